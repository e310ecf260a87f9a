#!/bin/bash
# step time of bench.py for a range of batch seeds (how much the per-rank batches of a multi-GPU run differ)
for sd in "$@"; do
  python bench.py --steps 8 --warmup 3 --seed $sd 2>/dev/null | tail -1 > /tmp/seed_$sd.json
  python - "$sd" <<'PY'
import json, sys
d = json.loads(open("/tmp/seed_%s.json" % sys.argv[1]).read())
print("seed", sys.argv[1], round(d["value"]), round(d["ms_per_step"], 3), d["config"]["mean_rounds"], round(d["config"]["alignment_gcups"]))
PY
done
