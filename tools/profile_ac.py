"""Driver for ncu launch lists of the ac_ppo iteration: python tools/profile_ac.py [num_envs] [iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppo_cpp_b200 import trainer  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 1
t = trainer.AcTrainer(env_id="HalfCheetah-v5", num_envs=n, num_steps=128, num_minibatches=4, update_epochs=4, seed=1)
t.reset(1)
for _ in range(iters):
    t.rollout(); t.gae(); t.update(0.0)
st = t.stats()
print(st["rollout_ms"], st["gae_ms"], st["update_ms"])
t.close()
