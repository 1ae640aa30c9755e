timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -2
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/r01_bench_v6.json 2> gpurun_out/r01_bench_v6.err; cat gpurun_out/r01_bench_v6.json; tail -3 gpurun_out/r01_bench_v6.err
