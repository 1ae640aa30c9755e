PB_IMBALANCE_PRINT=1 python tools/profile_walking.py Humanoid-v4 8192 tests/golden/walking_Humanoid-v4.npz 8 2 2>&1 | cut -c1-200
PB_ENV_ORDER=0 PB_IMBALANCE_PRINT=1 python tools/profile_walking.py Humanoid-v4 8192 tests/golden/walking_Humanoid-v4.npz 8 2 2>&1 | cut -c1-200 | sed 's/^/unordered: /'
