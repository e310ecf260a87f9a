for tag in yield_a spin_a yield_b spin_b; do
  mode=${tag%_*}
  SVP_SYNC=$mode python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 3 --c3-variants 0 > gpurun_out/n8_$tag.json 2> gpurun_out/n8_$tag.err
  python - $tag <<'PY'
import json, sys
d = json.loads([l for l in open("gpurun_out/n8_%s.json" % sys.argv[1]) if l.startswith("{")][-1])
print(sys.argv[1], "value %.0f e2e %.0f sync %.0f" % (d["value"], d["e2e"]["value"], d["e2e"]["sync_value"]))
PY
done
