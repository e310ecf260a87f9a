#!/bin/bash
# 8-GPU step time for different lane counts (host threads per rank) -- host-core oversubscription check
echo "host cores: $(nproc)"
for L in "$@"; do
  SVP_LANES=$L python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600+L)) bench.py --gpus 8 --steps 8 --warmup 3 2>/dev/null | tail -1 > /tmp/sl_$L.json
  python - "$L" <<'PY'
import json, sys
d = json.loads(open("/tmp/sl_%s.json" % sys.argv[1]).read())
print("lanes", sys.argv[1], round(d["value"]), round(d["ms_per_step"], 3), round(d["e2e"]["value"]))
PY
done
