# 1 GPU: validation at the new CTA-shape defaults: env / golden / edge / cli tests + config benches
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "env or golden or edge or cli" > gpurun_out/r01_pytest_cta.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01_pytest_cta.log
timeout 600 python tools/bench_configs.py 5 3 > gpurun_out/r01_bench_configs_v5.jsonl 2>/dev/null; cut -c1-300 gpurun_out/r01_bench_configs_v5.jsonl
