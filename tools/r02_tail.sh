#!/bin/bash
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -12) > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
timeout 200 python tools/ray_ab.py --phase setview --sets "4=0;4=1;4=2;4=3" --out set_ab.json 2>&1 | tail -4
timeout 200 python tools/ray_ab.py --phase integrate --sets "3=0" --out int_ab.json 2>&1 | tail -1
timeout 300 python bench.py --steps 40 --warmup 5 --no-cpu-baseline 2>gpurun_out/bench_tail.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench', d['ms_per_step'], d['phases'], d['e2e']['ms_per_step'], d['roofline']['frac'], d['gpu_launches'])"
tail -3 gpurun_out/bench_tail.err
