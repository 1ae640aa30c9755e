set -x
timeout 1200 python -m pytest tests -m gpu -q -k "env or golden" > gpurun_out/r01_pytest_gpu_v6.log 2>&1; tail -3 gpurun_out/r01_pytest_gpu_v6.log
for e in Humanoid-v4:8192 Ant-v5:16384 Hopper-v5:16384 HalfCheetah-v5:16384; do
  timeout 300 python tools/time_env_step.py ${e%%:*} ${e##*:} 60 20 2>&1 | tail -1
done | tee gpurun_out/r01_envstep_times_v6.log
