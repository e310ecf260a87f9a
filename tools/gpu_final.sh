# last validation of the round: GPU suite, then the bench line with the extras that fit the remaining budget
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r2f_pytest.log; cat gpurun_out/r2f_pytest.log
timeout 170 python bench.py --c3-variants 0 --no-c5-full > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err; tail -c 400 gpurun_out/r2_bench_final.json
