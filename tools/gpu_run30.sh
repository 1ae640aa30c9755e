# 1 GPU: CTA-shape variants for the Newton tasks and Humanoid 16-warp CTA, then env parity tests at the new defaults
mkdir -p gpurun_out
{
for t in 0 4 5; do echo "== HalfCheetah/Hopper PB_ENV_TUNING=$t"; PB_ENV_TUNING=$t python tools/bench_configs.py 5 3 C2,HC 2>/dev/null | cut -c1-290; done
for v in 0 11; do echo "== Humanoid PB_HUM_VARIANT=$v"; PB_HUM_VARIANT=$v python tools/bench_configs.py 5 3 C4/8 2>/dev/null | cut -c1-290; done
echo "== Ant default (16 warps)"; python tools/bench_configs.py 5 3 C3 2>/dev/null | cut -c1-290
} > gpurun_out/r01_cta_variants.txt 2>&1
cat gpurun_out/r01_cta_variants.txt
timeout 900 python -m pytest tests -m gpu -q -k "env or golden or edge" 2>&1 | tail -3
