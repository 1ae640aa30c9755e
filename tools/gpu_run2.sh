python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for v in 0 1 2 3; do PB_HUM_VARIANT=$v python tools/time_env_step.py Humanoid-v4 8192 25 20; done
