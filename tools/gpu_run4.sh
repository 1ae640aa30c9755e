# parity first, then timings of the env kernel for all tasks and a bench line
set -x
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r01_pytest_gpu_v4.log 2>&1; tail -15 gpurun_out/r01_pytest_gpu_v4.log
for e in Humanoid-v4:8192 Ant-v5:16384 Hopper-v5:4096 HalfCheetah-v5:8192; do
  timeout 300 python tools/time_env_step.py ${e%%:*} ${e##*:} 40 20 2>&1 | tail -1
done | tee gpurun_out/r01_envstep_times_v4.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_v4.json 2> gpurun_out/r01_bench_v4.err; cat gpurun_out/r01_bench_v4.json; tail -3 gpurun_out/r01_bench_v4.err
