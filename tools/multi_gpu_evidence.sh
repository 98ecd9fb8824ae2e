#!/bin/bash
# Multi-GPU evidence for one world size: exactness check on real GPUs, the bench line, the phase
# breakdown of the sharded ray cast.  Usage: gpurun --gpus N -- 'bash tools/multi_gpu_evidence.sh N'
N=$1
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 200 $RUN --master-port 29511 tools/sharded_check.py 2>&1 | grep -E "bit_identical|rror" | tail -2
timeout 300 $RUN --master-port 29512 bench.py --gpus $N --steps 40 --warmup 5 > gpurun_out/bench_ours_${N}gpu.json 2> gpurun_out/bench_ours_${N}gpu.err
python - $N <<'PY'
import json, sys
n = sys.argv[1]
d = json.loads(open("gpurun_out/bench_ours_%sgpu.json" % n).read().strip().splitlines()[-1])
print("bench", n, round(d["ms_per_step"], 4), {k: round(v, 4) for k, v in d["phases"].items() if k.endswith("_ms")},
      "e2e", round(d["e2e"]["ms_per_step"], 4), "value %.3g" % d["value"], d["clocks"])
PY
timeout 200 $RUN --master-port 29513 tools/sharded_phases.py 2>&1 | grep -E "\"world\"|rror" | tail -1
