python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r2f_pytest.log; cat gpurun_out/r2f_pytest.log
python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; tail -c 600 gpurun_out/r2_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference.json 2>/dev/null; tail -c 300 gpurun_out/r2_bench_reference.json
SVP_LANES=1 python bench.py --quick --steps 2 --warmup 1 > /dev/null 2>&1 && SVP_LANES=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2_launches.csv python bench.py --quick --steps 2 --warmup 1 > gpurun_out/r2_ncu.log 2>&1
