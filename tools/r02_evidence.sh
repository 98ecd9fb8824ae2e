#!/bin/bash
# Round-2 evidence in ONE GPU call (everything lands under gpurun_out/):
#   1 the GPU test suite                       4 ncu --set full of the REFERENCE's kernels K1-K14
#   2 both bench arms                          5 the reference's as-shipped build (no NDEBUG), timed once
#   3 launch list + ncu --set full of ours     6 compute-sanitizer memcheck / racecheck / synccheck logs
# Usage: gpurun --timeout 2400 -- 'bash tools/r02_evidence.sh [skip-tests]'
mkdir -p gpurun_out
if [ "$1" != "skip-tests" ]; then
  (timeout 1800 python -m pytest tests -m gpu -q -x 2>&1 | tail -15) > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
fi
timeout 300 python bench.py --impl reference --steps 40 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 300 python bench.py --steps 40 --warmup 5 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err
timeout 300 python bench.py --impl reference --ref-variant asshipped --steps 40 --warmup 5 --no-cpu-baseline \
    > gpurun_out/bench_ref_asshipped.json 2> gpurun_out/bench_ref_asshipped.err
python - <<'PY'
import json
for n in ("ours", "ref", "ref_asshipped"):
    try:
        d = json.loads(open("gpurun_out/bench_%s.json" % n).read().strip().splitlines()[-1])
        print(n, "ms/step %.4f" % d["ms_per_step"], {k: round(v, 4) for k, v in d["phases"].items() if k.endswith("_ms")},
              "e2e ms %.4f" % d["e2e"]["ms_per_step"], "roofline", d.get("roofline", {}).get("frac"))
    except Exception as e:
        print(n, "failed", e)
PY
# launch list of our arm, then one full capture of each hot kernel (only after the plain run exited 0)
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"vk_(raycast|integrate|alloc|bounds|carry|normals|setview)" \
    -s 18 -c 6 -o gpurun_out/prof_full -f python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/prof_full.ncu-rep --page raw --csv > gpurun_out/prof_full_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/prof_full_raw.csv > gpurun_out/ncu_full_summary.csv
# the reference's own kernels, same workload, same counters
python bench.py --impl reference --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/plain_ref.log 2>&1 &&
ncu --set full --clock-control none \
    -k regex:"UpdateBlockVisibility|CreateAllocationRequests|HandleAllocationRequests|IntegrateKernel|ComputePatches|ComputeBounds|ComputePoints|FillKernel|ComputeNormals|replace" \
    -s 45 -c 16 -o gpurun_out/prof_ref -f python bench.py --impl reference --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_ref.log 2>&1
ncu -i gpurun_out/prof_ref.ncu-rep --page raw --csv > gpurun_out/prof_ref_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/prof_ref_raw.csv > gpurun_out/ncu_ref_summary.csv
rm -f gpurun_out/prof_ref.ncu-rep
# sanitizers on config-1 frames (lock-free insert of the allocation kernel, shared-memory staging of the
# integrator and the ray marcher); the colliding-table variant forces every insert through the lock path
timeout 600 compute-sanitizer --tool memcheck python tools/one_frame.py --frames 2 > gpurun_out/sanitizer_memcheck_c1.log 2>&1
timeout 600 compute-sanitizer --tool memcheck python tools/one_frame.py --small --collide --frames 3 > gpurun_out/sanitizer_memcheck_collide.log 2>&1
timeout 900 compute-sanitizer --tool racecheck python tools/one_frame.py --small --frames 2 > gpurun_out/sanitizer_racecheck.log 2>&1
timeout 600 compute-sanitizer --tool synccheck python tools/one_frame.py --small --frames 2 > gpurun_out/sanitizer_synccheck.log 2>&1
tail -3 gpurun_out/sanitizer_*.log
# ray marcher A/B on the bench workload: distance mirror (knob 10) and look-ahead (knob 11)
timeout 300 python tools/ray_ab.py --sets "10=0,11=0;10=1,11=0;10=1,11=4;10=1,11=8;10=1,11=16;10=1,11=32" --out ray_ab_c2.json 2>&1 | tail -8
