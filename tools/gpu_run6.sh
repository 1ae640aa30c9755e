set -x
timeout 1200 python -m pytest tests -m gpu -q -k "env or golden or parity" > gpurun_out/r01_pytest_gpu_v5.log 2>&1; tail -5 gpurun_out/r01_pytest_gpu_v5.log
for o in 0 1; do
for e in Humanoid-v4:8192 Ant-v5:16384 Hopper-v5:16384 HalfCheetah-v5:16384; do
  PB_ENV_ORDER=$o timeout 300 python tools/time_env_step.py ${e%%:*} ${e##*:} 60 20 2>&1 | tail -1
done; done | tee gpurun_out/r01_envstep_times_v5.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_v5.json 2> gpurun_out/r01_bench_v5.err; cat gpurun_out/r01_bench_v5.json; tail -3 gpurun_out/r01_bench_v5.err
