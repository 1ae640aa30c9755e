mkdir -p gpurun_out
for w in 8 10; do
  PB_HUM_WPB=$w python tools/profile_walking.py Humanoid-v4 8192 tests/golden/walking_Humanoid-v4.npz 8 4 2>&1 | head -2 | cut -c1-170 | sed "s/^/WPB=$w /"
done | tee gpurun_out/r02_pgs_v5_ab.txt
timeout 900 python -m pytest tests/test_gpu_trained_state_parity.py tests/test_gpu_env_parity.py tests/test_golden.py -m gpu -q -x 2>&1 | tail -3
