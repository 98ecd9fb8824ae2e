#!/bin/bash
# One full ncu capture of one launch of every hot kernel of the bench step (run it only after
# tools/round_evidence.sh has passed on the same tree).  Usage: gpurun -- 'bash tools/round_ncu_full.sh'
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:"vk_(raycast|integrate|alloc|bounds|carry|normals)_kernel" \
    -s 18 -c 6 -o gpurun_out/prof_full -f python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
