"""One PPO iteration (Humanoid-v4, 8192 envs x 32 steps) with a short update, for ncu launch lists of the update kernels.
Usage: python tools/profile_update.py [epochs] [num_minibatches]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppo_cpp_b200 import trainer  # noqa: E402

ep = int(sys.argv[1]) if len(sys.argv) > 1 else 1
nmb = int(sys.argv[2]) if len(sys.argv) > 2 else 32
t = trainer.Trainer(env_id="Humanoid-v4", num_envs=8192, num_steps=32, num_minibatches=nmb, update_epochs=ep)
t.reset(1)
t.rollout(); t.gae()
t.update(3e-4)
t.update(3e-4)
st = t.stats()
print("update_ms", st["update_ms"], "per minibatch us", 1e3 * st["update_ms"] / (ep * nmb))
t.close()
