mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_v5.log 2>&1
tail -3 gpurun_out/pytest_gpu_v5.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_v2.json 2> gpurun_out/r02_bench_v2.err
tail -2 gpurun_out/r02_bench_v2.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_v2.json'))
print({k:d[k] for k in ('value','ms_per_step','phase_ms','rollout_only_sps')}, d['roofline']['kernel_ms'], d['e2e']['value'], d['cpu_baseline']['value'], d['e2e_env_step_host']['ms_per_step'])"
