set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r01_pytest_gpu.log
python bench.py --steps 3 --warmup 3 > gpurun_out/r01_bench_v1b.json 2> gpurun_out/r01_bench_v1b.err
python tools/profile_env_step.py Humanoid-v4 8192 30 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:env_step_kernel -s 28 -c 1 -f -o gpurun_out/r01_envstep_v1 python tools/profile_env_step.py Humanoid-v4 8192 30 > gpurun_out/ncu.log 2>&1
cat gpurun_out/r01_pytest_gpu.log gpurun_out/r01_bench_v1b.json gpurun_out/plain.log; tail -5 gpurun_out/ncu.log
