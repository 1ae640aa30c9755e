"""Generates the committed fixtures in tests/golden/ (run here, in the authoring container):

  randperm.npz        torch.randperm(n) on a CPU mt19937 generator for several (seed, n) -- the reference's minibatch
                      index source (src/ppo_continuous_action.cpp:490; src/ac_ppo_continuous_action.cpp:618,804).  torch in this
                      container IS an importable implementation of that operator, so these pin the permutation bit-exactly.
  ppo_step.npz        one clipped-PPO minibatch step (loss scalars, flat gradient, parameters after Adam) computed by torch
                      autograd on the CleanRL agent of src/ppo_continuous_action.cpp:120-169,496-541 for seeded inputs.
  env_step_<task>.npz single env-step transitions of the FP64 oracle from seeded states (reset(seed) obs, then one step
                      with seeded actions): regression pin for the oracle and a GPU-side check that needs no oracle run.
                      NOTE: the physics oracle itself is parity-unpinned w.r.t. MuJoCo 3.2.0.

Usage: python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import oracle_nn, pbo  # noqa: E402

ENVS = {"HalfCheetah-v5": (17, 6, 1.0), "Hopper-v5": (11, 3, 1.0), "Ant-v5": (105, 8, 1.0), "Humanoid-v4": (376, 17, 0.4)}


def randperm_fixture():
    out = {}
    for seed, n in [(1, 64), (1, 2048), (7, 2048), (1500, 4096), (1501, 1000), (12345, 65536)]:
        g = torch.Generator(device="cpu").manual_seed(seed)
        out[f"s{seed}_n{n}"] = torch.randperm(n, generator=g).numpy().astype(np.int64)
        out[f"s{seed}_n{n}_second"] = torch.randperm(n, generator=g).numpy().astype(np.int64)  # stream continues
    np.savez_compressed(os.path.join(HERE, "randperm.npz"), **out)


def ppo_step_fixture():
    torch.manual_seed(0)
    obs, act, B, mb = 11, 3, 256, 64
    agent = oracle_nn.Agent(obs, act)
    with torch.no_grad():
        agent.actor_logstd.normal_(0, 0.3)
        for p in agent.parameters():
            p.add_(0.05 * torch.randn_like(p))
    rng = np.random.default_rng(0)
    d = dict(obs=rng.normal(size=(B, obs)), actions=rng.normal(size=(B, act)) * 0.5, advantages=rng.normal(size=B), returns=rng.normal(size=B))
    d = {k: v.astype(np.float32) for k, v in d.items()}
    with torch.no_grad():
        _, lp, _, val = agent.get_action_and_value(torch.from_numpy(d["obs"]), torch.from_numpy(d["actions"]))
    d["logprobs"] = (lp.numpy() + rng.normal(size=B).astype(np.float32) * 0.15).astype(np.float32)
    d["values"] = (val.view(-1).numpy() + rng.normal(size=B).astype(np.float32) * 0.2).astype(np.float32)
    idx = rng.permutation(B)[:mb].astype(np.int64)
    out = dict(params0=agent.flat(), idx=idx, **d)
    opt = torch.optim.Adam(agent.parameters(), lr=3e-4, eps=1e-5)
    tb = {k: torch.from_numpy(v) for k, v in d.items()}
    ref = oracle_nn.ppo_minibatch_step(agent, opt, tb["obs"], tb["actions"], tb["logprobs"], tb["advantages"], tb["returns"], tb["values"],
                                       torch.from_numpy(idx))
    out["grad"] = ref["grad"]
    out["params1"] = agent.flat()
    out["scalars"] = np.array([ref[k] for k in ("policy_loss", "value_loss", "entropy", "old_approx_kl", "approx_kl", "clipfrac")], np.float64)
    np.savez_compressed(os.path.join(HERE, "ppo_step.npz"), **out)


def env_step_fixture():
    n = 16
    for env, (O, A, amax) in ENVS.items():
        v = pbo.VecEnv(env, n, chain=0)
        obs0 = v.reset(3).copy()
        rng = np.random.default_rng(5)
        acts = rng.uniform(-amax, amax, size=(3, n, A)).astype(np.float32)
        rec = dict(obs0=obs0, actions=acts)
        for t in range(3):
            o, r, te, tr, _ = v.step(acts[t])
            rec[f"obs{t + 1}"], rec[f"rew{t + 1}"], rec[f"term{t + 1}"], rec[f"trunc{t + 1}"] = o.copy(), r.copy(), te.copy(), tr.copy()
        np.savez_compressed(os.path.join(HERE, f"env_step_{env}.npz"), **rec)


if __name__ == "__main__":
    randperm_fixture()
    ppo_step_fixture()
    env_step_fixture()
    print("golden fixtures written to", HERE)
