# 1 GPU: round-1 final validation: full GPU suite, smoke, default bench (CPU baseline included), reference arm
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r01_pytest_gpu_v10.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01_pytest_gpu_v10.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r01_smoke_v10.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r01_smoke_v10.log
timeout 900 python bench.py > gpurun_out/r01_bench_v10.json 2> gpurun_out/r01_bench_v10.err; echo "bench rc=$?"; tail -2 gpurun_out/r01_bench_v10.err
grep -h '^{' gpurun_out/r01_bench_v10.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value']), d['phase_ms'], d['roofline_fp32']['frac'], round(d['e2e']['value']), d['cpu_baseline']['value'], d['clocks'])"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01_bench_ref_v10.json 2> gpurun_out/r01_bench_ref_v10.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/r01_bench_ref_v10.json
