# A/B: host wait modes (SVP_SYNC) on one GPU, configs[1]
run() { # label, env...
  label=$1; shift
  env "$@" python bench.py --quick --steps 20 --warmup 3 > gpurun_out/ab_$label.json 2>gpurun_out/ab_$label.err
  python - $label <<'PY'
import json, sys
d = json.load(open("gpurun_out/ab_%s.json" % sys.argv[1]))
print(sys.argv[1], "value %.0f e2e %.0f sync %.0f ms/step %.3f" % (d["value"], d["e2e"]["value"], d["e2e"]["sync_value"], d["ms_per_step"]))
PY
}
run yield SVP_SYNC=yield
run spin SVP_SYNC=spin
run yield2 SVP_SYNC=yield
run spin2 SVP_SYNC=spin
taskset -c 0-3 true && { run yield_4cores taskset -c 0-3 env SVP_SYNC=yield; run spin_4cores taskset -c 0-3 env SVP_SYNC=spin; run yield_4cores2 taskset -c 0-3 env SVP_SYNC=yield; run spin_4cores2 taskset -c 0-3 env SVP_SYNC=spin; }
