set -x
timeout 900 python -m pytest tests -m gpu -x -q -k "ppo or golden or cli or trainer" > gpurun_out/r01_pytest_dw1b.log 2>&1; tail -3 gpurun_out/r01_pytest_dw1b.log
PB_WG_OVERLAP=0 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_dw1b_serial.json 2> gpurun_out/r01_bench_dw1b_serial.err; tail -2 gpurun_out/r01_bench_dw1b_serial.err
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_dw1b_overlap.json 2> gpurun_out/r01_bench_dw1b_overlap.err; tail -2 gpurun_out/r01_bench_dw1b_overlap.err
PB_DW1_TC=0 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_dw1b_off.json 2> gpurun_out/r01_bench_dw1b_off.err; tail -2 gpurun_out/r01_bench_dw1b_off.err
for f in serial overlap off; do grep -h '^{' gpurun_out/r01_bench_dw1b_$f.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$f', round(d['value']), d['phase_ms'], d['gpu_launches'])"; done
