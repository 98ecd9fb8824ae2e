#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/ray_ab.py --sets "11=0;11=4;11=8;11=12;11=8,7=16;11=6,7=12" --out ray_ab_c2.json 2>&1 | tail -6
VK_TUNE="11=8" timeout 200 python tools/tune_bench.py --what timeline 2>&1 | tail -40 | head -16
timeout 400 python -m pytest tests/test_extract_gpu.py tests/test_parity_gpu.py -m gpu -q -x -k "oracle_exact or watertight or c1_full" 2>&1 | tail -4
