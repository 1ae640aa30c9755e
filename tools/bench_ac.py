"""C5 workload (BASELINE.json configs[4]): HalfCheetah-v5 ac_ppo_continuous_action, 8192 envs per GPU, T=128,
4 minibatches x 4 epochs (the ac defaults, src/ac_ppo_continuous_action.cpp:66-67).  Single GPU phase times and SPS of the
AcTrainer path; one JSON line.  usage: python tools/bench_ac.py [env_id] [num_envs] [steps] [warmup]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppo_cpp_b200 import trainer  # noqa: E402

env_id = sys.argv[1] if len(sys.argv) > 1 else "HalfCheetah-v5"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
warmup = int(sys.argv[4]) if len(sys.argv) > 4 else 2
T = 128
t = trainer.AcTrainer(env_id=env_id, num_envs=N, num_steps=T, num_minibatches=4, update_epochs=4)
t.reset(1)
for _ in range(warmup):
    t.iteration(2.5e-4)
t.sync()
t0 = time.perf_counter()
ph = []
for _ in range(steps):
    t.iteration(2.5e-4)
    st = t.stats()
    ph.append((st["rollout_ms"], st["gae_ms"], st["update_ms"]))
dt = time.perf_counter() - t0
print(json.dumps({"workload": f"{env_id} ac_ppo: {N} envs x {T} steps, 4 minibatches x 4 epochs", "value": N * T * steps / dt,
                  "unit": "env-steps/s", "ms_per_iteration": dt / steps * 1e3,
                  "phase_ms": {"rollout": ph[-1][0], "gae": ph[-1][1], "update": ph[-1][2]}, "n_gpus": 1}))
t.close()
