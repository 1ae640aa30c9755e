mkdir -p gpurun_out
for t in 0 2 1; do
  PB_ENV_TUNING=$t python tools/profile_walking.py Humanoid-v4 8192 tests/golden/walking_Humanoid-v4.npz 8 4 2>&1 | head -1 | sed "s/^/TUNING=$t /"
done | tee gpurun_out/r02_tuning_ab.txt
