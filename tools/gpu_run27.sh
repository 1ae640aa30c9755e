# 1 GPU: register-resident Newton factor/solve: env parity + golden + edge tests, then config benches
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "env or golden or edge or cli" > gpurun_out/r01_pytest_newton.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r01_pytest_newton.log
timeout 600 python tools/bench_configs.py 5 3 > gpurun_out/r01_bench_configs_v2.jsonl 2> gpurun_out/r01_bench_configs_v2.err; cut -c1-330 gpurun_out/r01_bench_configs_v2.jsonl; tail -3 gpurun_out/r01_bench_configs_v2.err
