for v in 100 625; do for m in wide narrow; do SVP_MIX_REGULAR=$m python bench.py --quick --variants $v --steps 10 --warmup 3 > gpurun_out/ab_${m}_$v.json 2>/dev/null; done; done
python - <<'PY'
import json
for v in (100, 625):
    for m in ("wide", "narrow"):
        d = json.load(open("gpurun_out/ab_%s_%d.json" % (m, v))); r = d["roofline"]; k = d["kernel_ms_per_step"]
        print(v, m, "value %.0f e2e %.0f frac %.3f mixed_ms %.3f cells_ratio %.3f fails %.1f single-part %s" % (d["value"], d["e2e"]["value"], r["frac"], k["k1_score_mixed"], r["score_pass_cells_computed_over_algorithmic"], r["mixed_rounds"]["certificates_failed_inside_a_mixed_launch_per_step"], r["timed"][-20:]))
PY
