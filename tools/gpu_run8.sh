timeout 600 python -m pytest tests -m gpu -q -k "env or golden" 2>&1 | tail -2
for v in 0 14 13; do PB_HUM_VARIANT=$v timeout 300 python tools/time_env_step.py Humanoid-v4 8192 60 20 2>&1 | tail -1; done
for t in 0 2 3; do for e in Ant-v5:16384 Hopper-v5:16384 HalfCheetah-v5:16384; do
  PB_ENV_TUNING=$t timeout 300 python tools/time_env_step.py ${e%%:*} ${e##*:} 60 20 2>&1 | tail -1 | sed "s/^/tuning=$t /"
done; done
