mkdir -p gpurun_out
for w in 8 9 10; do
  PB_HUM_WPB=$w python tools/profile_walking.py Humanoid-v4 8192 tests/golden/walking_Humanoid-v4.npz 8 4 2>&1 | head -1 | sed "s/^/WPB=$w /"
done | tee gpurun_out/r02_wpb_ab.txt
