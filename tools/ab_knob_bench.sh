#!/bin/bash
# A/B of one VK_TUNE setting on the full bench line (device-resident step and end-to-end), alternating runs.
# usage: tools/ab_knob_bench.sh "<VK_TUNE for B>" [runs] [steps]     (A = defaults)
B=$1; RUNS=${2:-3}; STEPS=${3:-20}
mkdir -p gpurun_out
for i in $(seq 1 $RUNS); do
  timeout 120 python bench.py --steps $STEPS --warmup 5 --no-cpu-baseline > gpurun_out/ab_A_$i.json
  VK_TUNE="$B" timeout 120 python bench.py --steps $STEPS --warmup 5 --no-cpu-baseline > gpurun_out/ab_B_$i.json
done
python - "$B" $RUNS <<'PY'
import json, sys
import numpy as np
b, runs = sys.argv[1], int(sys.argv[2])
out = {}
for arm in "AB":
    rows = [json.loads(open("gpurun_out/ab_%s_%d.json" % (arm, i)).read().strip().splitlines()[-1]) for i in range(1, runs + 1)]
    out[arm] = dict(knobs=("defaults" if arm == "A" else b), step_ms=[r["ms_per_step"] for r in rows],
                    e2e_ms=[r["e2e"]["ms_per_step"] for r in rows], sync_every_step_ms=[r["e2e"]["sync_every_step"]["ms_per_step"] for r in rows],
                    phases_ms=[{k: round(v, 4) for k, v in r["phases"].items() if k.endswith("_ms")} for r in rows],
                    host_submit_us=[r["e2e"]["host_submit_us_median"] for r in rows])
    print(arm, out[arm]["knobs"], "step", np.round(out[arm]["step_ms"], 4), "e2e", np.round(out[arm]["e2e_ms"], 4),
          "sync", np.round(out[arm]["sync_every_step_ms"], 4), "submit", out[arm]["host_submit_us"])
json.dump(out, open("gpurun_out/ab_knob_bench.json", "w"), indent=1)
PY
