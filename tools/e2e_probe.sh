#!/bin/bash
# where the end-to-end leg's time goes: the streamed pipeline under diagnostics settings
# usage: tools/e2e_probe.sh "<VK_TUNE a>;<VK_TUNE b>;..." "<VULCAN_PIPELINE_DEBUG values>"
IFS=';' read -ra TUNES <<< "${1:-15=1;15=0}"
for dbg in ${2:-0 1 2 3}; do
  for tune in "${TUNES[@]}"; do
    VULCAN_PIPELINE_DEBUG=$dbg VK_TUNE="$tune" timeout 120 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/e2e_probe.json
    python - $dbg "$tune" <<'PY'
import json, sys
d = json.loads(open("gpurun_out/e2e_probe.json").read().strip().splitlines()[-1])
print("debug", sys.argv[1], "tune", sys.argv[2], "step %.4f e2e %.4f sync %.4f submit %.1f" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["sync_every_step"]["ms_per_step"], d["e2e"]["host_submit_us_median"]))
PY
  done
done
