PB_IMBALANCE_PRINT=1 python tools/profile_walking.py Humanoid-v4 8192 tests/golden/walking_Humanoid-v4.npz 8 2 2>&1 | cut -c1-200 | grep -E "walking|imbalance|solver per"
timeout 900 python -m pytest tests/test_gpu_trained_state_parity.py tests/test_gpu_env_parity.py tests/test_golden.py -m gpu -q -x 2>&1 | tail -3
