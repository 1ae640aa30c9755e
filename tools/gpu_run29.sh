# 1 GPU: PGS one-shuffle sweep: env parity/golden/edge tests + Humanoid bench (twice)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "env or golden or edge" > gpurun_out/r01_pytest_pgs.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r01_pytest_pgs.log
for i in 1 2; do
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), d['phase_ms'], round(d['roofline']['kernel_ms'],3), round(d['e2e']['value']))"
done
