cd /root/repo
timeout 600 python -m pytest tests/test_gpu_env_parity.py tests/test_gpu_trained_state_parity.py -x -q -m gpu 2>&1 | tail -3
for B in 1 0; do
PB_ENV_BALANCE=$B timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('hum PB_ENV_BALANCE=$B', round(d['value']), d['phase_ms'], d['roofline']['kernel_ms'])"
PB_ENV_BALANCE=$B timeout 300 python bench.py --workload ac --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ac PB_ENV_BALANCE=$B', round(d['value']), d['phase_ms'])"
PB_ENV_TUNING=3 PB_ENV_BALANCE=$B timeout 300 python bench.py --workload ac --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ac 2x12 PB_ENV_BALANCE=$B', round(d['value']), d['phase_ms'])"
done
PB_AC_GEMM_DEBUG=1 timeout 120 python tools/profile_ac.py 8192 1 2>&1 | grep gemm256 | sed -n '1,2p;9,10p'
