"""SURVEY 8d workloads C2 / C3 (and C4 at one GPU) through the Trainer path: PPO iteration SPS and phase times, at the pinned
"walking" state of each task (tests/golden/walking_<env>.npz: policy + normaliser after 50 iterations, lr = 0, at least 8 untimed
rollouts - the protocol of bench.py).  One JSON line per config.  usage: python tools/bench_configs.py [steps] [warmup] [configs]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppo_cpp_b200 import trainer  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
warmup = int(sys.argv[2]) if len(sys.argv) > 2 else 3
only = sys.argv[3].split(",") if len(sys.argv) > 3 else None
for name, env_id, N in [("C2", "Hopper-v5", 4096), ("C3", "Ant-v5", 16384), ("C4/8", "Humanoid-v4", 8192), ("HC", "HalfCheetah-v5", 8192)]:
    if only and name not in only:
        continue
    T = 32
    t = trainer.Trainer(env_id=env_id, num_envs=N, num_steps=T, num_minibatches=32, update_epochs=10, seed=1)
    t.reset(1)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    t.load_snapshot(os.path.join(root, "tests", "golden", f"walking_{env_id}.npz"))
    for _ in range(max(0, 8 - warmup)):
        t.rollout()
    for _ in range(warmup):
        t.iteration(0.0)
    t.sync()
    t0 = time.perf_counter()
    for _ in range(steps):
        t.iteration(0.0)
    t.sync()
    dt = time.perf_counter() - t0
    st = t.stats()
    print(json.dumps({"config": name, "workload": f"{env_id} PPO at the pinned walking state (lr = 0): {N} envs x {T} steps, 32 minibatches x 10 epochs, 1 GPU",
                      "value": N * T * steps / dt, "unit": "env-steps/s", "ms_per_iteration": dt / steps * 1e3,
                      "phase_ms_last": {"rollout": st["rollout_ms"], "gae": st["gae_ms"], "update": st["update_ms"]},
                      "episodes_last_iter": st["episodes"], "overflow_envs": st["overflow_envs"]}), flush=True)
    t.close()
