#!/bin/bash
# bench A/B over several VK_TUNE settings: tools/ab_multi.sh "set1;set2;..." [runs]
IFS=';' read -ra TUNES <<< "$1"
RUNS=${2:-2}
for i in $(seq 1 $RUNS); do
  for tune in "${TUNES[@]}"; do
    VK_TUNE="$tune" timeout 120 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/abm.json
    python - "$tune" <<'PY'
import json, sys
d = json.loads(open("gpurun_out/abm.json").read().strip().splitlines()[-1])
print("tune [%s] step %.4f e2e %.4f phases %s" % (sys.argv[1], d["ms_per_step"], d["e2e"]["ms_per_step"], {k: round(v, 4) for k, v in d["phases"].items() if k.endswith("_ms")}))
PY
  done
done
