#!/bin/bash
# Runs on an 8-GPU box (gpurun --gpus 8 -- bash tools/gpu_scale.sh [N ...]): bench.py at N = 1, 2, 4, 8 (or the listed N) the way
# the driver launches it.
NS=${@:-1 2 4 8}
for n in $NS; do
  if [ $n = 1 ]; then
    python bench.py --gpus 1 --steps 20 --warmup 3 --c4-variants 0 --no-c5-full --whole-program-records 0 > gpurun_out/r2_scale_n1.json 2> gpurun_out/r2_scale_n1.err
    continue
  fi
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/r2_scale_n$n.json 2> gpurun_out/r2_scale_n$n.err
  tail -c 300 gpurun_out/r2_scale_n$n.err
done
python - <<'PY'
import json
for n in (1, 2, 4, 8):
    try:
        d = json.loads([l for l in open("gpurun_out/r2_scale_n%d.json" % n) if l.startswith("{")][-1])
        c = d.get("sharded_candidate_list") or {}
        print(n, "weak value %.0f e2e %.0f | sharded C3: e2e %.0f resident %.0f imbalance %.3f gather %.3f s pinned %s" % (
            d["value"], d["e2e"]["value"], c.get("e2e_variants_per_s", 0), c.get("resident_variants_per_s", 0),
            c.get("rank_imbalance_max_over_mean", 0), c.get("gather_seconds_rank0", 0), d["config"]["host_cpus_pinned_per_rank"]))
    except Exception as e:
        print(n, "failed", e)
PY
