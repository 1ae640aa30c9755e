"""Measured sensitivity of one env step to each MuJoCo behaviour the oracle restates from memory (oracle/ASSUMPTIONS.md).

For every task: the 256 harvested walking states (tests/golden/walking_<env>.npz) are stepped once (frame_skip x mj_step) with
the recorded actions, with the restatement (variant 0) and with one alternative at a time (oracle/pbo_physics.c pbo_variant);
reported: median and maximum over the states of the relative change max_i |x_i' - x_i| / max(1, |x_i|) of qpos and qvel, next to
the fp32-vs-fp64 agreement the GPU parity tests require (1e-3 on 99 % of the contact-active states).
Usage: python tools/oracle_sensitivity.py > profiles/r02_oracle_sensitivity.txt"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import pbo  # noqa: E402

VARIANTS = [(1, "warm start kept fixed through RK4 stages 1-3"), (2, "contacts / rows in reverse pair order"),
            (4, "exact diagonal of J M^-1 J^T instead of diagApprox"), (8, "default contact frame for plane-capsule"),
            (16, "solver tolerance x 100"), (32, "solver iterations x 2"), (64, "explicit joint damping in Euler")]
FRAME_SKIP = {"HalfCheetah-v5": 5, "Hopper-v5": 4, "Ant-v5": 5, "Humanoid-v4": 5}
ACT = {"HalfCheetah-v5": 1.0, "Hopper-v5": 1.0, "Ant-v5": 1.0, "Humanoid-v4": 0.4}
L = pbo.lib()
L.pbo_set_variant.argtypes = [C.c_int]


def step_all(env, d, bits):
    L.pbo_set_variant(bits)
    s = pbo.Sim(env)
    out = []
    for i in range(len(d["qpos"])):
        s.reset()
        s.qpos[:] = d["qpos"][i]; s.qvel[:] = d["qvel"][i]; s.qacc_warmstart[:] = d["warm"][i]
        s.ctrl[:] = np.clip(d["act"][i], -ACT[env], ACT[env])
        s.step(FRAME_SKIP[env])
        out.append(np.concatenate([s.qpos, s.qvel]))
    L.pbo_set_variant(0)
    return np.array(out)


for env in ("Humanoid-v4", "Ant-v5", "Hopper-v5", "HalfCheetah-v5"):
    d = np.load(os.path.join("tests", "golden", f"walking_{env}.npz"))
    nq = d["qpos"].shape[1]
    base = step_all(env, d, 0)
    print(f"== {env}: one env step from {len(base)} walking states")
    for bits, name in VARIANTS:
        alt = step_all(env, d, bits)
        rel = np.abs(alt - base) / np.maximum(1.0, np.abs(base))
        eq, ev = rel[:, :nq].max(axis=1), rel[:, nq:].max(axis=1)
        print(f"   {name:52s} qpos median {np.median(eq):.1e} max {eq.max():.1e} | qvel median {np.median(ev):.1e} max {ev.max():.1e}")
