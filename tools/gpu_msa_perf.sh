#!/bin/bash
# GPU box: build_consensus with the fills on the device vs SeqAn's own globalMsaAlignment, at the crop sizes of configs[1] (k = 400:
# ~800 bp) and configs[3] (k = 1000 + kb insertions); plus the K1 roofline at larger batches.
python -m pytest tests/test_gpu_msa.py -m gpu -x -q 2>&1 | tail -3
for L in 800 3000 7000; do
  n=8; [ $L -ge 3000 ] && n=2; [ $L -ge 7000 ] && n=1
  python tools/msa_perf_cases.py /tmp/msa_$L.txt $L $n
  echo "== crops of $L bp, $n case(s): device fills"; ./sviper_b200/host/msa_harness /tmp/msa_$L.txt 2 > /tmp/dev_$L.out
  if [ $L -le 3000 ]; then echo "== SeqAn on the host"; SVIPER_B200_MSA=host ./sviper_b200/host/msa_harness /tmp/msa_$L.txt 2 > /tmp/host_$L.out; cmp /tmp/dev_$L.out /tmp/host_$L.out && echo "consensus identical"; fi
done 2>&1 | tee gpurun_out/r2_msa_perf.log
for v in 625 2000; do python bench.py --quick --variants $v --steps 6 --warmup 3 > gpurun_out/r2_bench_v$v.json 2>/dev/null; done
python - <<'PY'
import json
for v in (625, 2000):
    d = json.load(open("gpurun_out/r2_bench_v%d.json" % v)); r = d["roofline"]
    print(v, "value %.0f e2e %.0f k1 all-launch frac %.3f full-wave %.3f whole-step-alg %.3f k5 share %.3f" % (d["value"], d["e2e"]["value"], r["frac"], r["full_wave"]["frac"], r["whole_step_algorithmic"]["frac"], d["final_alignment_kernel"]["share_of_step"]))
PY
