# A/B of the certificate prediction over batch sizes: configs[1] at 100 / 625 variants per step, configs[3]
run() { # label, variants, c4, env...
  label=$1; nv=$2; c4=$3; shift 3
  env "$@" python bench.py --variants $nv --c3-variants 0 --c4-variants $c4 --no-c5-full --whole-program-records 0 --steps 10 --warmup 3 > gpurun_out/ab_$label.json 2>/dev/null
  python - $label <<'PY'
import json, sys
d = json.load(open("gpurun_out/ab_%s.json" % sys.argv[1])); r = d["roofline"]; k = d["kernel_ms_per_step"]; x = d.get("large_insertions") or {}
print(sys.argv[1], "value %.0f e2e %.0f frac %.3f mixed_ms %.3f cells %.3f fails %.1f | C4 %s var/s" % (d["value"], d["e2e"]["value"], r["frac"], k["k1_score_mixed"], r["score_pass_cells_computed_over_algorithmic"], r["mixed_rounds"]["certificates_failed_inside_a_mixed_launch_per_step"], x.get("variants_per_s")))
PY
}
run p100 100 200 A=1
run n100 100 200 SVP_NO_PREDICT=1
run p625 625 0 A=1
