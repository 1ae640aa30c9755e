cd /root/repo
timeout 900 python -m pytest tests/test_gpu_env_parity.py tests/test_gpu_trained_state_parity.py tests/test_gpu_edge_cases.py tests/test_cli.py -x -q -m gpu 2>&1 | tail -2
for A in 1 0; do
PB_ENV_ORDER_ASYNC=$A timeout 300 python bench.py --steps 10 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ppo ASYNC=$A', round(d['value']), d['phase_ms'])"
PB_ENV_ORDER_ASYNC=$A timeout 300 python bench.py --workload ac --steps 5 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ac ASYNC=$A', round(d['value']), d['phase_ms'])"
done
