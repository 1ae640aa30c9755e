set -x
python tools/profile_env_step.py Humanoid-v4 8192 60 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:env_step_kernel -s 58 -c 1 -f -o gpurun_out/r01_envstep_v5 python tools/profile_env_step.py Humanoid-v4 8192 60 > gpurun_out/ncu.log 2>&1
cat gpurun_out/plain.log; tail -3 gpurun_out/ncu.log
