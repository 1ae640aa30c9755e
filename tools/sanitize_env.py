"""Small run of every env kernel for compute-sanitizer (memcheck / racecheck): 64 envs x 12 steps per task through the host API,
plus two PPO iterations of a tiny Humanoid trainer (policy, GAE, update kernels incl. the tcgen05 ones)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppo_cpp_b200 import capi, trainer  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 12
rng = np.random.default_rng(0)
for env in ("HalfCheetah-v5", "Hopper-v5", "Ant-v5", "Humanoid-v4"):
    v = capi.ParVectorEnv(env, 64, chain=capi.PB_CHAIN_PPO)
    v.reset(1)
    for _ in range(steps):
        v.step(rng.normal(size=(64, v.act_dim)).astype(np.float32))
    print(env, "ok", v.counters())
    v.close()
if len(sys.argv) > 2:
    t = trainer.Trainer(env_id="Humanoid-v4", num_envs=128, num_steps=4, num_minibatches=2, update_epochs=1)
    t.reset(1)
    t.iteration(3e-4)
    print("trainer ok", t.stats()["value_loss"])
    t.close()
