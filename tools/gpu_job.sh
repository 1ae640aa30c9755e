timeout 900 python -m pytest tests/test_cli.py tests/test_gpu_ac_parity.py -m gpu -q -x 2>&1 | tail -15
