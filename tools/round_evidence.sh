#!/bin/bash
# One GPU call that produces the round's evidence under gpurun_out/: GPU test log, both bench arms,
# and the ncu launch list.  Usage: gpurun -- 'bash tools/round_evidence.sh'; the full ncu capture of the
# hot kernels is a second call: gpurun -- 'bash tools/round_ncu_full.sh'
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6) > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 40 --warmup 5 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err
timeout 300 python bench.py --impl reference --steps 40 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python - <<'PY'
import json
for n in ("ours", "ref"):
    try:
        d = json.load(open("gpurun_out/bench_%s.json" % n))
        print(n, "ms/step %.4f" % d["ms_per_step"], {k: round(v, 4) for k, v in d["phases"].items() if k.endswith("_ms")},
              "e2e ms %.4f" % d["e2e"]["ms_per_step"], "roofline", d.get("roofline", {}).get("frac"))
    except Exception as e:
        print(n, "failed", e)
PY
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
# pyramid + depth tracker: ours against the reference build, closed-loop drift, launch list,
# one full capture of the iteration kernel
timeout 300 python tools/track_bench.py > gpurun_out/track_bench.json 2> gpurun_out/track_bench.err
timeout 300 python tools/tracked_orbit.py --frames 36 > gpurun_out/tracked_orbit.json 2> gpurun_out/tracked_orbit.err
python tools/track_once.py > gpurun_out/track_once.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/track_launches.csv \
    python tools/track_once.py > gpurun_out/track_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"vk_track_iter_kernel|vk_downsample_frame_kernel" -s 4 -c 3 \
    -o gpurun_out/prof_track -f python tools/track_once.py > gpurun_out/track_ncu_full.log 2>&1
ncu -i gpurun_out/prof_track.ncu-rep --page raw --csv > gpurun_out/track_ncu_full_raw.csv 2>/dev/null
tail -2 gpurun_out/track_ncu_full.log
