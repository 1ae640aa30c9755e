set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:env_step_kernel --launch-skip 25 -c 1 -o gpurun_out/r01_envstep_ant python tools/profile_env_step.py Ant-v5 16384 30 > gpurun_out/ncu_ant.log 2>&1
tail -3 gpurun_out/ncu_ant.log | cut -c1-200
