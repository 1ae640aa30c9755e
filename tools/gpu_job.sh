cd /root/repo
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/r02_pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02_pytest_gpu_final.log
