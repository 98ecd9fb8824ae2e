#!/bin/bash
# per-warp timeline of the ray marcher under different knobs + the GPU tests the first run did not reach
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "replicated or sharded_volume" 2>&1 | tail -15) > gpurun_out/pytest_gpu_shard.log; cat gpurun_out/pytest_gpu_shard.log
(timeout 900 python -m pytest tests/test_stepwise_gpu.py tests/test_stress_gpu.py tests/test_tracker_gpu.py tests/test_snapshot_gpu.py -m gpu -q 2>&1 | tail -15) > gpurun_out/pytest_gpu_rest.log; cat gpurun_out/pytest_gpu_rest.log
for k in "11=0" "11=8" "11=0,1=1,2=0" "11=8,1=1,2=0"; do
  echo "== knobs $k"
  VK_TUNE="$k" timeout 200 python tools/tune_bench.py --what timeline 2>&1 | tail -40
done > gpurun_out/timeline_knobs.log 2>&1
head -120 gpurun_out/timeline_knobs.log
