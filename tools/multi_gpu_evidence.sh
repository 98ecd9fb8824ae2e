#!/bin/bash
# Multi-GPU evidence for one world size: exactness + per-phase timing on real GPUs for the given configs,
# then the bench line.  Usage: gpurun --gpus N -- 'bash tools/multi_gpu_evidence.sh N "small c2 c4 c5"'
N=$1; CONFIGS=${2:-"small c2"}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29511
for c in $CONFIGS; do
  timeout 600 $RUN --master-port $port tools/sharded_check.py --config $c > gpurun_out/sharded_check_${N}gpu_${c}.log 2>&1
  grep -E "bit_identical" gpurun_out/sharded_check_${N}gpu_${c}.log | tail -1 | cut -c1-900 || tail -5 gpurun_out/sharded_check_${N}gpu_${c}.log
  grep -E "rror|Traceback" gpurun_out/sharded_check_${N}gpu_${c}.log | head -5
  port=$((port+1))
done
timeout 300 $RUN --master-port 29531 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_ours_${N}gpu.json 2> gpurun_out/bench_ours_${N}gpu.err
python - $N <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/bench_ours_%sgpu.json" % n).read().strip().splitlines()[-1])
    print("bench", n, round(d["ms_per_step"], 4), {k: round(v, 4) for k, v in d["phases"].items() if k.endswith("_ms")},
          "e2e", round(d["e2e"]["ms_per_step"], 4), "value %.3g" % d["value"], d["clocks"])
except Exception as e:
    print("bench failed", e)
    print(open("gpurun_out/bench_ours_%sgpu.err" % n).read()[-1500:])
PY
