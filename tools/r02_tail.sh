#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/shard_emul.py 2 interleaved 2>&1 | tail -20
timeout 120 python tools/shard_emul.py 2 phased 2>&1 | tail -20
for k in "11=0" "11=8" "11=0,1=1,2=0" "11=8,1=1,2=0"; do
  echo "== knobs $k"
  VK_TUNE="$k" timeout 200 python tools/tune_bench.py --what timeline 2>&1 | tail -40
done > gpurun_out/timeline_knobs.log 2>&1
head -150 gpurun_out/timeline_knobs.log
