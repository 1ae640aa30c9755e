rm -rf models/pytest_*
timeout 900 python -m pytest tests/test_cli.py -m gpu -q -x 2>&1 | tail -15
