#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_cxx_dropin.py tests/test_pipeline_gpu.py -m gpu -q -x 2>&1 | tail -8
timeout 300 python bench.py --steps 40 --warmup 5 2>gpurun_out/bench_tail.err > gpurun_out/bench_tail.json; python -c "
import json
d=json.loads(open('gpurun_out/bench_tail.json').read().strip().splitlines()[-1]); print('bench', d['ms_per_step'], d['phases'], 'e2e', d['e2e']['ms_per_step'], d['e2e']['host_submit_us_median'], d['e2e']['sync_every_step'], d['roofline']['frac'], d['gpu_launches'])"
tail -3 gpurun_out/bench_tail.err
