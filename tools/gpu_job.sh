cd /root/repo
timeout 900 python -m pytest tests/test_gpu_ac_parity.py -x -q -m gpu > gpurun_out/pytest_ac_fused.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_ac_fused.log
timeout 300 python bench.py --workload ac --steps 5 --warmup 2 --no-cpu-baseline > gpurun_out/r02_bench_ac_fused2.json 2> gpurun_out/r02_bench_ac_fused2.err; echo "bench rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/r02_bench_ac_fused2.json').read().strip().splitlines()[-1]); print(d['value'], d['phase_ms'])"; tail -3 gpurun_out/r02_bench_ac_fused2.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_launches_ac_fused2.csv python bench.py --workload ac --steps 2 --warmup 2 --no-cpu-baseline --profile-iteration > gpurun_out/ncu_launches_ac.log 2>&1
wc -l gpurun_out/r02_launches_ac_fused2.csv
