# 1 GPU: Humanoid env kernel CTA-shape variants at the bench state (wave quantisation: 8192 envs over resident warps)
mkdir -p gpurun_out
{
for v in 0 9 8 13 2 1 12; do
  echo "== PB_HUM_VARIANT=$v"
  PB_HUM_VARIANT=$v timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), d['phase_ms'], round(d['roofline']['kernel_ms'],3), round(d['e2e']['value']))"
done
} > gpurun_out/r01_hum_variants_v8.txt 2>&1
cat gpurun_out/r01_hum_variants_v8.txt
