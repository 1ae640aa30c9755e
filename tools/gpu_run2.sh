python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for tn in 0 1; do
PB_ENV_TUNING=$tn python tools/time_env_step.py Ant-v5 16384 25 20
PB_ENV_TUNING=$tn python tools/time_env_step.py Hopper-v5 4096 25 20
PB_ENV_TUNING=$tn python tools/time_env_step.py HalfCheetah-v5 8192 25 20
done
