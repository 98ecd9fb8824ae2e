#!/bin/bash
# A/B of two builds of the library on the bench step: tools/ab_bench.sh vulcan_b200/old.so [rounds]
# (the alternative .so is swapped in on the GPU box's scratch copy only)
alt=$1; rounds=${2:-2}
cp vulcan_b200/libvulcan_b200.so /tmp/vk_new.so
for i in $(seq $rounds); do
  for v in new alt; do
    if [ $v = alt ]; then cp $alt vulcan_b200/libvulcan_b200.so; else cp /tmp/vk_new.so vulcan_b200/libvulcan_b200.so; fi
    timeout 200 python bench.py --steps 40 --warmup 5 --no-cpu-baseline 2>/dev/null > /tmp/ab.json
    python - $v <<'PY'
import json, sys
d = json.loads(open("/tmp/ab.json").read().strip().splitlines()[-1])
print(sys.argv[1], round(d["ms_per_step"], 4), {k: round(v, 4) for k, v in d["phases"].items() if k.endswith("_ms")}, "e2e", round(d["e2e"]["ms_per_step"], 4))
PY
  done
done
cp /tmp/vk_new.so vulcan_b200/libvulcan_b200.so
