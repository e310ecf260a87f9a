#!/bin/bash
# Runs on the GPU box (gpurun -- bash tools/gpu_checks.sh): full GPU suite, fuzz campaign with its report, ncu launch list of
# the bench command (single part), compute-sanitizer memcheck + racecheck over the small parity tests. Outputs in gpurun_out/.
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2f_pytest.log; cat gpurun_out/r2f_pytest.log
python tools/fuzz_parity.py 300 500000 gpurun_out/r2_fuzz.json > gpurun_out/r2_fuzz.log 2>&1; tail -2 gpurun_out/r2_fuzz.log
SVP_LANES=1 python bench.py --quick --steps 2 --warmup 1 > /dev/null 2>&1 && SVP_LANES=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2_launches.csv python bench.py --quick --steps 2 --warmup 1 > gpurun_out/r2_ncu.log 2>&1
tail -2 gpurun_out/r2_ncu.log | cut -c1-200
timeout 900 compute-sanitizer --tool memcheck --log-file gpurun_out/r2_memcheck.log python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2_memcheck_pytest.log 2>&1
tail -3 gpurun_out/r2_memcheck_pytest.log; tail -3 gpurun_out/r2_memcheck.log
timeout 1200 compute-sanitizer --tool racecheck --log-file gpurun_out/r2_racecheck.log python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "polish or final or round or align" > gpurun_out/r2_racecheck_pytest.log 2>&1
tail -3 gpurun_out/r2_racecheck_pytest.log; tail -3 gpurun_out/r2_racecheck.log
