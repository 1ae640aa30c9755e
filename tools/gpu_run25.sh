# 1 GPU: edge-case tests, config benches (C2, C3, C4/8, HalfCheetah), full suite
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_edge_cases.py -q > gpurun_out/r01_pytest_edge.log 2>&1; echo "edge rc=$?"; tail -25 gpurun_out/r01_pytest_edge.log
timeout 600 python tools/bench_configs.py 5 3 > gpurun_out/r01_bench_configs.jsonl 2> gpurun_out/r01_bench_configs.err; cat gpurun_out/r01_bench_configs.jsonl; tail -3 gpurun_out/r01_bench_configs.err
