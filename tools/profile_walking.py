"""Driver for timing / ncu at the pinned "walking" state: python tools/profile_walking.py [env_id] [num_envs] [npz] [warm] [timed]
Loads the policy snapshot (tools/make_walking_checkpoint.py), runs `warm` rollouts with the frozen policy so episodes and
contacts reach their stationary mix, then `timed` rollouts and prints the rollout / per-step time.
ncu: `-k regex:env_step_kernel --launch-skip $((32*warm+8)) -c 1` lands inside the first timed rollout."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppo_cpp_b200 import trainer  # noqa: E402

env_id = sys.argv[1] if len(sys.argv) > 1 else "Humanoid-v4"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
npz = sys.argv[3] if len(sys.argv) > 3 else f"tests/golden/walking_{env_id}.npz"
warm = int(sys.argv[4]) if len(sys.argv) > 4 else 3
timed = int(sys.argv[5]) if len(sys.argv) > 5 else 3
t = trainer.Trainer(env_id=env_id, num_envs=n, num_steps=32, num_minibatches=32, update_epochs=10, seed=1)
t.reset(1)
t.load_snapshot(npz)
for _ in range(warm):
    t.rollout()
ms = []
for _ in range(timed):
    t.rollout()
    st = t.stats()
    ms.append(st["rollout_ms"])
print(f"{env_id} N={n} walking state: rollout ms {['%.2f' % m for m in ms]} -> {sum(ms) / len(ms) / 32:.3f} ms/env-step-launch (incl. policy), "
      f"episodes/rollout {st['episodes']}, mean return {st['episodic_return_sum'] / max(1, st['episodes']):.1f}, "
      f"mean length {st['episodic_length_sum'] / max(1, st['episodes']):.1f}, overflow {st['overflow_envs']}, "
      f"variant={os.environ.get('PB_HUM_VARIANT', '0')}")
import ctypes as C
from ppo_cpp_b200 import capi
c2 = (C.c_ulonglong * 2)()
capi.lib().pb_env_solver_counters(capi.lib().pb_trainer_env(t._h), c2)
print(f"solver: {c2[0] / max(1, st['forward_evals']):.2f} PGS sweeps and {c2[1] / max(1, st['forward_evals']):.1f} row steps per forward evaluation "
      f"(oracle at the harvested states: 20.0 / 149)")
t.close()
