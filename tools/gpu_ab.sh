# A/B: the shard of configs[2] through the asynchronous scheduler vs one blocking call (5,000 and 625 records on one GPU)
run() { # label, c3 records, env...
  label=$1; n=$2; shift 2
  env "$@" python bench.py --c3-variants $n --c4-variants 0 --no-c5-full --whole-program-records 0 --steps 5 --warmup 3 > gpurun_out/ab_$label.json 2>gpurun_out/ab_$label.err
  python - $label <<'PY'
import json, sys
d = json.load(open("gpurun_out/ab_%s.json" % sys.argv[1])); c = d.get("sharded_candidate_list") or {}
print(sys.argv[1], "value %.0f e2e %.0f | C3 e2e %.0f (%.1f ms, gather %.1f ms) resident %.0f" % (d["value"], d["e2e"]["value"], c.get("e2e_variants_per_s", 0), 1e3 * c.get("e2e_seconds", 0), 1e3 * c.get("gather_seconds_rank0", 0), c.get("resident_variants_per_s", 0)))
PY
}
run async5000 5000 A=1
run block5000 5000 BENCH_C3_BLOCKING=1
run async625 625 A=1
run block625 625 BENCH_C3_BLOCKING=1
