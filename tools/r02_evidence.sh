#!/bin/bash
# Round-2 evidence in ONE GPU call (everything lands under gpurun_out/):
#   1 the GPU test suite                       4 ncu --set full of the REFERENCE's kernels K1-K14
#   2 both bench arms                          5 the reference's as-shipped build (no NDEBUG), timed once
#   3 launch list + ncu --set full of ours     6 bulk-copy integrator capture, knob A/Bs, config / extract / track benches
# (compute-sanitizer is closed on this pool: bounds are checked by the stress tests' own asserts instead)
# Usage: gpurun --timeout 2400 -- 'bash tools/r02_evidence.sh [skip-tests]'
mkdir -p gpurun_out
if [ "$1" != "skip-tests" ]; then
  (timeout 1800 python -m pytest tests -m gpu -q -x 2>&1 | tail -15) > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
fi
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err
timeout 300 python bench.py --impl reference --ref-variant asshipped --steps 20 --warmup 5 --no-cpu-baseline \
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
ncu --set full --clock-control none --import-source on -k regex:"vk_(raycast|integrate|alloc|bounds|carry|normals)_kernel" \
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
# the bulk-copy (cp.async.bulk + mbarrier) integrator variant, for the record of the A/B (knob 14)
VK_TUNE="14=1" ncu --set full --clock-control none -k regex:"vk_integrate_bulk" -s 3 -c 2 -o gpurun_out/prof_bulk -f \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bulk.log 2>&1
ncu -i gpurun_out/prof_bulk.ncu-rep --page raw --csv > gpurun_out/prof_bulk_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/prof_bulk_raw.csv > gpurun_out/ncu_bulk_summary.csv
rm -f gpurun_out/prof_bulk.ncu-rep gpurun_out/prof_bulk_raw.csv
# knob A/Bs on the bench workload: distance mirror (10), paired free-space steps (11), L2 warm-up (12), bulk integrator (14)
timeout 300 python tools/ray_ab.py --sets "10=1;10=0;11=0;12=0" --out ray_ab_c2.json 2>&1 | tail -6
timeout 300 python tools/ray_ab.py --phase integrate --sets "14=0;14=1" --out integrate_ab_c2.json 2>&1 | tail -4
# the other BASELINE configurations per phase against the reference, surface extraction, tracker
timeout 900 python tools/config_bench.py > gpurun_out/config_bench.log 2> gpurun_out/config_bench.err; tail -c 2500 gpurun_out/config_bench.log
timeout 300 python tools/extract_bench.py > gpurun_out/extract_bench.json 2> gpurun_out/extract_bench.err; tail -c 800 gpurun_out/extract_bench.json
timeout 300 python tools/track_bench.py > gpurun_out/track_bench.json 2> gpurun_out/track_bench.err; tail -c 800 gpurun_out/track_bench.json
# the extraction kernels (points + faces) of the final tree, one full capture each
ncu --set full --clock-control none -k regex:"vk_extract" -s 2 -c 3 -o gpurun_out/prof_extract -f \
    python tools/extract_bench.py > gpurun_out/ncu_extract.log 2>&1
ncu -i gpurun_out/prof_extract.ncu-rep --page raw --csv > gpurun_out/prof_extract_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/prof_extract_raw.csv > gpurun_out/ncu_extract_summary.csv
rm -f gpurun_out/prof_extract.ncu-rep gpurun_out/prof_extract_raw.csv
