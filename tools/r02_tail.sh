#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_extract_gpu.py tests/test_cxx_dropin.py -m gpu -q -x 2>&1 | tail -15
timeout 200 python tools/extract_bench.py > gpurun_out/extract_bench.json 2> gpurun_out/extract_bench.err; tail -3 gpurun_out/extract_bench.json | cut -c1-600
