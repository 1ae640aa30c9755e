set -x
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -3
python tools/profile_env_step.py Humanoid-v4 8192 60 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:env_step_kernel -s 58 -c 1 -f -o gpurun_out/r01_envstep_v6 python tools/profile_env_step.py Humanoid-v4 8192 60 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
