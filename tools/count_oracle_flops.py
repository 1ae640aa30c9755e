"""Algorithmic FLOPs of one env step, COUNTED by running the FP64 oracle with an operation-counting scalar type
(oracle/count/count.h: every add, mul, div, sqrt and transcendental the restatement executes) over the benchmark's states: the 256
harvested walking states of each task, one env step (frame_skip x mj_step + rnePostConstraint where the task reads cfrc_ext) with
the recorded actions.  SURVEY.md 8d: this is the numerator of roofline_fp32 ("achieved" = counted FLOPs x envs / kernel time).
Usage: python tools/count_oracle_flops.py  -> profiles/r02_oracle_flops.json"""
import ctypes as C
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
lib = os.path.join(ROOT, "oracle", "count", "liboracle_count.so")
if not os.path.exists(lib):
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "count"])
os.environ["PBO_LIB"] = lib
import numpy as np  # noqa: E402

from oracle import pbo  # noqa: E402

L = pbo.lib()
L.pbo_count_get.argtypes = [C.POINTER(C.c_uint64)]
FRAME_SKIP = {"HalfCheetah-v5": 5, "Hopper-v5": 4, "Ant-v5": 5, "Humanoid-v4": 5}
ACT = {"HalfCheetah-v5": 1.0, "Hopper-v5": 1.0, "Ant-v5": 1.0, "Humanoid-v4": 0.4}
CFRC = {"Ant-v5", "Humanoid-v4"}
out = {}
for env in ("Humanoid-v4", "Ant-v5", "Hopper-v5", "HalfCheetah-v5"):
    d = np.load(os.path.join(ROOT, "tests", "golden", f"walking_{env}.npz"))
    s = pbo.Sim(env)
    tot, evals, nefc, nit = np.zeros(6), 0, [], []
    for i in range(len(d["qpos"])):
        s.reset()
        s.qpos[:] = d["qpos"][i]; s.qvel[:] = d["qvel"][i]; s.qacc_warmstart[:] = d["warm"][i]
        s.ctrl[:] = np.clip(d["act"][i], -ACT[env], ACT[env])
        n0 = s.counts()["nforward"]
        s.trace_reset()
        L.pbo_count_reset()
        s.step(FRAME_SKIP[env])
        if env in CFRC:
            s.rne_post()
        c = (C.c_uint64 * 6)()
        L.pbo_count_get(c)
        tot += np.array(list(c), dtype=np.float64)
        evals += s.counts()["nforward"] - n0
        a, b = s.trace()
        nefc += list(a); nit += list(b)
    n = len(d["qpos"])
    add, mul, div, sq, tr, cmp_ = tot / n
    flop = add + mul + div + sq + tr
    out[env] = {"flop_per_env_step": flop, "flop_per_forward_evaluation": flop * n / evals, "add": add, "mul": mul, "div": div, "sqrt": sq,
                "transcendental": tr, "compare": cmp_, "forward_evaluations_per_env_step": evals / n, "mean_nefc": float(np.mean(nefc)),
                "mean_solver_iterations": float(np.mean(nit)), "states": n}
    print(f"{env}: {flop / 1e6:.3f} MFLOP per env step ({flop * n / evals / 1e3:.1f} k per forward evaluation; add {add:.0f} mul {mul:.0f} div {div:.0f} "
          f"sqrt {sq:.0f} trans {tr:.0f}); mean nefc {np.mean(nefc):.1f}, solver iterations {np.mean(nit):.1f}")
json.dump({"source": "tools/count_oracle_flops.py: FP64 oracle with an op-counting scalar over the 256 walking states of each task, one env step each",
           "tasks": out}, open(os.path.join(ROOT, "profiles", "r02_oracle_flops.json"), "w"), indent=1)
