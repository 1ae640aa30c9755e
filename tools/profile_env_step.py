"""Small driver for ncu: N Humanoid-v4 (or other) envs, reset, then STEPS env steps with random actions through the
host API.  Usage: python tools/profile_env_step.py [env_id] [num_envs] [steps]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppo_cpp_b200 import capi  # noqa: E402

env_id = sys.argv[1] if len(sys.argv) > 1 else "Humanoid-v4"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 30
v = capi.ParVectorEnv(env_id, n, chain=capi.PB_CHAIN_PPO)
v.reset(1)
rng = np.random.default_rng(0)
hi = v.get_action_space_max()
t0 = time.perf_counter()
for s in range(steps):
    a = rng.normal(0.0, 1.0, size=(n, v.act_dim)).astype(np.float32)
    o, r, te, tr, _ = v.step(a)
dt = time.perf_counter() - t0
print(f"{env_id} N={n} steps={steps}: {dt / steps * 1e3:.2f} ms/step host API, done frac last step {float((te + tr).mean()):.3f}, counters {v.counters()}")
v.close()
