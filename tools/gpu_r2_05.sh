set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1
tail -5 gpurun_out/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_v4.log 2>&1
tail -5 gpurun_out/pytest_gpu_v4.log
