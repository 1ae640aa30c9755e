"""Train each task for ITERS PPO iterations on the GPU (through the trainer C ABI) and write the "walking" fixture
tests/golden/walking_<env>.npz:

  params            flat fp32 policy/critic parameters after ITERS iterations (torch registration order)
  ob_mean/ob_var    NormalizeObservation statistics averaged over the envs, ob_count their median count, rw4 the reward
                    normaliser's (mean, var, acc, count) averaged - the policy is only meaningful with its normaliser
  qpos/qvel/warm    float64 states of K envs harvested in the middle of a rollout under the trained policy (contacts
                    active), act = the policy's sampled action of the same env one step earlier (same distribution)

The fixture pins two things: the parity tests step CUDA and oracle from these states
(tests/test_gpu_trained_state_parity.py), and bench.py times its iterations from this frozen policy (lr = 0), so the
dominant kernel sees the same contact load in bench.py, under ncu and in the driver's run.

Run on a GPU box: python tools/make_walking_checkpoint.py [out_dir] [iters]  (writes to gpurun_out/ by default; copy the
files to tests/golden/ afterwards).
"""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppo_cpp_b200 import capi, trainer  # noqa: E402

OUT = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
ITERS = int(sys.argv[2]) if len(sys.argv) > 2 else 50
K = 256
TASKS = [("Humanoid-v4", 8192), ("Ant-v5", 16384), ("Hopper-v5", 4096), ("HalfCheetah-v5", 8192)]
if len(sys.argv) > 3:
    TASKS = [t for t in TASKS if t[0] in sys.argv[3:]]
T = 32
os.makedirs(OUT, exist_ok=True)
L = capi.lib()
fp = capi._fp

for env_id, N in TASKS:
    t = trainer.Trainer(env_id=env_id, num_envs=N, num_steps=T, num_minibatches=32, update_epochs=10, seed=1)
    t.reset(1)
    t0 = time.perf_counter()
    hist = []
    for it in range(ITERS):
        t.iteration(3e-4)
        st = t.stats()
        hist.append((st["rollout_ms"], st["update_ms"], st["episodic_return_sum"] / max(1, st["episodes"]), st["episodes"]))
    print(f"{env_id}: {ITERS} iterations x {N} envs x {T} steps in {time.perf_counter() - t0:.1f} s; "
          f"rollout ms first/last {hist[0][0]:.1f}/{hist[-1][0]:.1f}, mean episodic return first/last {hist[0][2]:.1f}/{hist[-1][2]:.1f}, "
          f"overflow_envs {st['overflow_envs']}", flush=True)
    params = t.get_params()
    eh = L.pb_trainer_env(t._h)
    O, A = L.pb_env_obs_dim(eh), L.pb_env_act_dim(eh)
    nq, nv, nu = C.c_int(), C.c_int(), C.c_int()
    L.pb_env_model_dims(eh, C.byref(nq), C.byref(nv), C.byref(nu))
    nq, nv = nq.value, nv.value
    # normaliser statistics: average over a sample of envs
    ms, vs, cs, rws = [], [], [], []
    for i in range(0, N, max(1, N // 128)):
        m, v, c, rw = np.zeros(O, np.float32), np.zeros(O, np.float32), C.c_float(), np.zeros(4, np.float32)
        capi._check(L.pb_env_get_norm_stats(eh, i, m.ctypes.data_as(fp), v.ctypes.data_as(fp), C.byref(c), rw.ctypes.data_as(fp)))
        ms.append(m); vs.append(v); cs.append(c.value); rws.append(rw)
    ob_mean, ob_var = np.mean(ms, axis=0).astype(np.float32), np.mean(vs, axis=0).astype(np.float32)
    ob_count, rw4 = float(np.median(cs)), np.mean(rws, axis=0).astype(np.float32)
    # harvest: one more rollout under the trained policy, states at its end
    t.rollout()
    t.sync()
    acts = t.read("actions").reshape(T, N, A)[T - 1]
    done = (t.read("term") + t.read("trunc")) > 0
    idx = [i for i in range(N) if not done[i]][:: max(1, N // (2 * K))][:K]
    qpos, qvel, warm = np.zeros((len(idx), nq)), np.zeros((len(idx), nv)), np.zeros((len(idx), nv))
    dp = capi._dp
    for k, i in enumerate(idx):
        capi._check(L.pb_env_get_state(eh, i, qpos[k].ctypes.data_as(dp), qvel[k].ctypes.data_as(dp), warm[k].ctypes.data_as(dp)))
    path = os.path.join(OUT, f"walking_{env_id}.npz")
    np.savez_compressed(path, params=params, ob_mean=ob_mean, ob_var=ob_var, ob_count=np.float32(ob_count), rw4=rw4,
                        qpos=qpos, qvel=qvel, warm=warm, act=acts[idx].astype(np.float32), env_index=np.array(idx),
                        iters=ITERS, num_envs=N, num_steps=T, seed=1,
                        history=np.array(hist, np.float32))
    print(f"  wrote {path}: {len(idx)} states, z range {qpos[:, 1 if nq < 12 else 2].min():.3f}..{qpos[:, 1 if nq < 12 else 2].max():.3f}, "
          f"ob_count {ob_count:.1f}", flush=True)
    t.close()
