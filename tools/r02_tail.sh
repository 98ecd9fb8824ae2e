#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/ray_ab.py --phase setview --sets "4=0;4=9" --out set_ab.json 2>&1 | tail -2
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_stress_gpu.py -m gpu -q -x 2>&1 | tail -3
timeout 300 python bench.py --steps 40 --warmup 5 --no-cpu-baseline 2>gpurun_out/bench_tail.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench', d['ms_per_step'], d['phases'], d['e2e']['ms_per_step'], d['roofline']['frac'], d['gpu_launches'])"
