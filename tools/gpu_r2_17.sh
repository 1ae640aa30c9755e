timeout 900 python -m pytest tests/test_gpu_edge_cases.py tests/test_gpu_ppo_parity.py tests/test_gpu_ac_parity.py -m gpu -q -x 2>&1 | tail -12
