#!/bin/bash
# the sharded end-to-end leg with copies switched off (diagnostics): tools/e2e_probe_sharded.sh N
N=${1:-2}
for dbg in 0 1 3 0; do
  VULCAN_PIPELINE_DEBUG=$dbg timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 \
      bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/e2e_probe_sh.json 2> gpurun_out/e2e_probe_sh.err
  python - $dbg <<'PY'
import json, sys
d = json.loads(open("gpurun_out/e2e_probe_sh.json").read().strip().splitlines()[-1])
print("debug", sys.argv[1], "step %.4f e2e %.4f sync %.4f submit %.1f" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["sync_every_step"]["ms_per_step"], d["e2e"]["host_submit_us_median"]))
PY
done
