#!/bin/bash
# GPU box: the round's profile evidence. (1) bench line without a profiler, (2) ncu launch list of the same command with the
# batch as one part, (3) one ncu --set full capture each of a mixed-round score launch and of a local MSA fill.
set -x
python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_reference.json 2>/dev/null
SVP_LANES=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 2600 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --quick --steps 2 --warmup 1 > gpurun_out/r2_ncu.log 2>&1
SVP_LANES=1 ncu --set full --clock-control none --import-source on -k regex:k1_score_kernel --launch-skip 2 -c 1 -f -o gpurun_out/r2_k1_mixed \
    python bench.py --quick --steps 1 --warmup 1 > gpurun_out/r2_ncu_k1.log 2>&1
python tools/msa_perf_cases.py /tmp/msa_3000.txt 3000 1
ncu --set full --clock-control none --import-source on -k regex:msa_fill_kernel --launch-skip 1 -c 1 -f -o gpurun_out/r2_k6_msa_local \
    ./sviper_b200/host/msa_harness /tmp/msa_3000.txt > gpurun_out/r2_ncu_k6.log 2>&1
ls -la gpurun_out/*.ncu-rep
