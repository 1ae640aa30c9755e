mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ac_parity.py tests/test_gpu_ppo_parity.py tests/test_gpu_edge_cases.py -m gpu -q -x 2>&1 | tail -8
python bench.py --workload ac --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_ac_tc2.json 2> gpurun_out/r02_bench_ac_tc2.err
tail -2 gpurun_out/r02_bench_ac_tc2.err
python -c "
import json; d=json.loads(open('gpurun_out/r02_bench_ac_tc2.json').read().strip().splitlines()[-1]); print(d['value'], d['phase_ms'], d['roofline']['achieved'])"
