cd /root/repo
timeout 900 python -m pytest tests/test_gpu_ppo_parity.py tests/test_gpu_edge_cases.py tests/test_cli.py -x -q -m gpu 2>&1 | tail -3
for P in 1 0; do
PB_POLICY_TC=$P timeout 300 python bench.py --steps 10 --warmup 5 --no-cpu-baseline 2>gpurun_out/pol$P.err | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('PB_POLICY_TC=$P', round(d['value']), d['phase_ms'], d['gpu_launches'])"
done
tail -2 gpurun_out/pol1.err
