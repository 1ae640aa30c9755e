"""Parity + timing report of the CUDA vector env against the CPU oracle (run on the GPU box).

Prints, per task: reset-observation error, single env-step state error from identical states (split into steps
with and without active constraints), multi-step rollout drift, overflow counters, and a first timing of the step
kernel.  Used to calibrate the tolerances written into tests/test_gpu_env_parity.py.
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import pbo  # noqa: E402
from ppo_cpp_b200 import capi  # noqa: E402

ENVS = ["HalfCheetah-v5", "Hopper-v5", "Ant-v5", "Humanoid-v4"]


def relerr(a, b):
    return np.abs(a - b) / np.maximum(1.0, np.abs(b))


def main():
    np.set_printoptions(precision=3, linewidth=200)
    N = 64
    for env in ENVS:
        print("=" * 20, env)
        ov = pbo.VecEnv(env, N, chain=0)
        gv = capi.ParVectorEnv(env, N, chain=0)
        o_obs = ov.reset(1).copy()
        g_obs = gv.reset(1).copy()
        print("reset obs max relerr", relerr(g_obs, o_obs).max())
        # single step from identical states
        rng = np.random.default_rng(0)
        sims = [ov.sim(i) for i in range(N)]
        for lift in (0.0, 1.0):
            ov.reset(1)
            gv.reset(1)
            zi = 1 if env in ("HalfCheetah-v5", "Hopper-v5") else 2
            for i, s in enumerate(sims):
                s.qpos[zi] += lift
                s.qvel[:] = s.qvel + rng.normal(size=s.nv) * 0.5
                s.qacc_warmstart[:] = 0
                s.forward()  # normalises the quaternion, sets qacc_warmstart = qacc
                gv.set_state(i, s.qpos.copy(), s.qvel.copy(), np.zeros(s.nv))
            gv.forward()
            act = rng.uniform(-1, 1, size=(N, ov.act_dim)).astype(np.float32)
            oo = [x.copy() for x in ov.step(act)[:4]]
            go = [x.copy() for x in gv.step(act)[:4]]
            eq, ev = [], []
            nefc = []
            for i, s in enumerate(sims):
                q, v, w = gv.get_state(i)
                eq.append(relerr(q, s.qpos).max())
                ev.append(relerr(v, s.qvel).max())
                nefc.append(s.counts()["nefc"])
            eq, ev, nefc = np.array(eq), np.array(ev), np.array(nefc)
            print(f"lift={lift}: qpos relerr max {eq.max():.2e} med {np.median(eq):.2e} | qvel max {ev.max():.2e} med "
                  f"{np.median(ev):.2e} | nefc(last eval) mean {nefc.mean():.1f} max {nefc.max()}")
            print("   obs relerr max", relerr(go[0], oo[0]).max(), "rew abs err max", np.abs(go[1] - oo[1]).max(),
                  "term mismatch", int((go[2] != oo[2]).sum()))
        # rollout drift
        ov.reset(1)
        gv.reset(1)
        for t in range(40):
            act = rng.normal(size=(N, ov.act_dim)).astype(np.float32)
            oo = ov.step(act)
            go = gv.step(act)
            if t % 5 == 4 or t < 3:
                e = relerr(go[0], oo[0]).max(axis=1)
                print(f" t={t:2d} obs relerr med {np.median(e):.2e} p90 {np.percentile(e, 90):.2e} max {e.max():.2e} | rew err "
                      f"{np.abs(go[1] - oo[1]).max():.2e} | done mismatch {(go[2] != oo[2]).sum() + (go[3] != oo[3]).sum()}")
        print("counters (forwards, overflows):", gv.counters())
        # wrapper chain parity on identical raw streams is covered by the ppo chain run below
        ov2 = pbo.VecEnv(env, N, chain=1)
        gv2 = capi.ParVectorEnv(env, N, chain=1)
        e0 = relerr(gv2.reset(1), ov2.reset(1)).max()
        errs = []
        for t in range(10):
            act = rng.normal(size=(N, ov.act_dim)).astype(np.float32)
            oo = ov2.step(act)
            go = gv2.step(act)
            errs.append((relerr(go[0], oo[0]).max(), np.abs(go[1] - oo[1]).max()))
        print("ppo chain: reset err", e0, "step errs (obs, rew)", errs[0], errs[4], errs[9])
        gv.close()
        gv2.close()
        # timing
        for n in (4096, 8192, 16384):
            g = capi.ParVectorEnv(env, n, chain=1)
            g.reset(1)
            act = rng.normal(size=(n, ov.act_dim)).astype(np.float32)
            for _ in range(3):
                g.step(act)
            t0 = time.time()
            K = 10
            for _ in range(K):
                g.step(act)
            dt = (time.time() - t0) / K
            print(f"  N={n}: host-API step {dt * 1e3:.2f} ms -> {n / dt:.3e} env-steps/s (incl. H2D/D2H)")
            g.close()


if __name__ == "__main__":
    main()
