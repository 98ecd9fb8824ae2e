#!/bin/bash
mkdir -p gpurun_out
VK_TUNE="14=1" timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "oracle_exact or c1_full or baseline_configs" 2>&1 | tail -4
timeout 200 python tools/ray_ab.py --phase integrate --sets "14=0;14=1;14=1,3=4;14=1,3=5;14=1,3=6;14=1,3=7" --out int_ab_c2.json 2>&1 | tail -6
timeout 300 python tools/ray_ab.py --phase integrate --scene room --frames 4 --sets "14=0;14=1;14=1,3=6;14=1,3=7" --out int_ab_room.json 2>&1 | tail -4
