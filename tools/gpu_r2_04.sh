set -x
mkdir -p gpurun_out
python tools/profile_walking.py Humanoid-v4 8192 tests/golden/walking_Humanoid-v4.npz 6 4 > gpurun_out/walk_time_hum_v2.log 2>&1
cat gpurun_out/walk_time_hum_v2.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_v3.log 2>&1
tail -25 gpurun_out/pytest_gpu_v3.log
