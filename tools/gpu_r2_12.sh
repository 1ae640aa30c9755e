mkdir -p gpurun_out
for n in 1184 2368 4736 8192; do
  python tools/profile_walking.py Humanoid-v4 $n tests/golden/walking_Humanoid-v4.npz 8 4 2>&1 | head -1 | cut -c1-150
done | tee gpurun_out/r02_alone_vs_shared.txt
PB_ENV_TUNING=2 python tools/profile_walking.py Humanoid-v4 8192 tests/golden/walking_Humanoid-v4.npz 8 4 2>&1 | head -1 | cut -c1-150 | sed "s/^/1x16: /" | tee -a gpurun_out/r02_alone_vs_shared.txt
