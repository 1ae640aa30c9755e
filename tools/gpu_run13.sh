set -x
timeout 900 python -m pytest tests -m gpu -x -q -k "ppo or golden or cli or trainer" > gpurun_out/r01_pytest_dw1.log 2>&1; tail -5 gpurun_out/r01_pytest_dw1.log
PB_DW1_TC=0 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_dw1_off.json 2> gpurun_out/r01_bench_dw1_off.err; tail -2 gpurun_out/r01_bench_dw1_off.err
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_dw1_on.json 2> gpurun_out/r01_bench_dw1_on.err; tail -2 gpurun_out/r01_bench_dw1_on.err
grep -h '^{' gpurun_out/r01_bench_dw1_off.json gpurun_out/r01_bench_dw1_on.json | cut -c1-400
