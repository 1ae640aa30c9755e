set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ac_parity.py tests/test_cli.py -m gpu -q > gpurun_out/r01_pytest_ac128.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01_pytest_ac128.log
for m in 1 0; do echo "PB_AC_GEMM128=$m"; PB_AC_GEMM128=$m timeout 300 python tools/bench_ac.py HalfCheetah-v5 8192 3 2 2>/dev/null; done | tee gpurun_out/r01_bench_ac_gemm128.txt
