"""Two-GPU ac_ppo_continuous_action update (BASELINE.json configs[4]; src/ac_ppo_continuous_action.cpp:816-888) through the C ABI:
every rank owns different rows; the advantage normalisation uses the CROSS-RANK moments of :830-849 (allreduce-avg of the local
means, then allreduce-sum of sum (A - mean)^2 with Bessel's correction over world x minibatch_per_device), the flat gradient is
averaged over the ranks (:878-885) - all three exchanges through the LL peer-memory kernel (csrc/pb_peer.cuh).  Checked against torch autograd run per rank with those moments and averaged on the host:
gradient rel 2e-4, parameters 2e-5 after the Adam step, replicas bit-identical.  Needs two devices; skipped otherwise."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV, OBS, ACT, N, T, NMB = "HalfCheetah-v5", 17, 6, 64, 8, 4

WORKER = r"""
import os, sys, time
import numpy as np
sys.path.insert(0, sys.argv[1])
rank, world, td = int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
import torch
from oracle import oracle_nn_ac
from ppo_cpp_b200 import trainer
env, obs, act, N, T, nmb = "HalfCheetah-v5", 17, 6, 64, 8, 4
B = N * T
t = trainer.AcTrainer(env_id=env, num_envs=N, num_steps=T, num_minibatches=nmb, update_epochs=1, device=rank, rank=rank, world_size=world, seed=3)
p0 = t.get_params()
agent = oracle_nn_ac.AcAgent(obs, act, 1.0, -1.0)
agent.load_flat(p0)
rng = np.random.default_rng(100 + rank)            # every rank owns different rows
data = dict(obs=rng.normal(size=(B, obs)), actions=rng.uniform(-1, 1, size=(B, act)) * 0.98, advantages=rng.normal(size=B) * (1.0 + rank) + 0.3 * rank,
            returns=rng.normal(size=B))
data = {k: v.astype(np.float32) for k, v in data.items()}
with torch.no_grad():
    _, lp, _, val = agent.get_action_and_value(torch.from_numpy(data["obs"]), torch.from_numpy(data["actions"]))
data["logprobs"] = (lp.numpy() + rng.normal(size=B).astype(np.float32) * 0.08).astype(np.float32)
data["values"] = (val.view(-1).numpy() + rng.normal(size=B).astype(np.float32) * 0.2).astype(np.float32)
for k, v in data.items():
    t.write(k, v)
uid_path = os.path.join(td, "uid")
if rank == 0:
    uid = trainer.comm_unique_id()
    with open(uid_path + ".tmp", "wb") as f:
        f.write(uid)
    os.rename(uid_path + ".tmp", uid_path)
else:
    for _ in range(600):
        if os.path.exists(uid_path):
            break
        time.sleep(0.1)
    uid = open(uid_path, "rb").read()
t.comm_init(uid, rank, world)
from ppo_cpp_b200 import capi
assert capi.lib().pb_ac_trainer_exchange_mode(t._h) == 2   # the LL peer-memory exchange, not the NCCL fallback
t.broadcast_params()
idx = np.random.default_rng(7 + rank).permutation(B)[: B // nmb].astype(np.int32)   # each rank draws its own permutation (:618)
t.update_minibatch(idx, 2.5e-4)
np.savez(os.path.join(td, f"r{rank}.npz"), p0=p0, grad=t.read("grad"), params=t.get_params(), idx=idx, **data)
t.close()
"""


@pytest.mark.timeout(600)
def test_ac_update_two_gpus_cross_rank_moments_and_gradient_average():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from oracle import oracle_nn_ac
    world = 2
    with tempfile.TemporaryDirectory() as td:
        script = os.path.join(td, "w.py")
        with open(script, "w") as f:
            f.write(WORKER)
        procs = [subprocess.Popen([sys.executable, script, ROOT, str(r), str(world), td], stderr=subprocess.PIPE, text=True) for r in range(world)]
        errs = []
        for p in procs:
            try:
                _, e = p.communicate(timeout=300)
            except subprocess.TimeoutExpired:
                for q in procs:
                    q.kill()
                raise
            errs.append(e)
        assert all(p.returncode == 0 for p in procs), errs
        res = [dict(np.load(os.path.join(td, f"r{r}.npz"))) for r in range(world)]
    assert np.array_equal(res[0]["p0"], res[1]["p0"])
    assert np.array_equal(res[0]["grad"], res[1]["grad"]) and np.array_equal(res[0]["params"], res[1]["params"])   # replicas identical
    # cross-rank advantage moments (:830-849)
    mb = [r["advantages"][r["idx"]].astype(np.float64) for r in res]
    mean = np.mean([m.mean() for m in mb])
    var = sum(((m - mean) ** 2).sum() for m in mb) / (world * len(mb[0]) - 1)
    moments = (torch.tensor(np.float32(mean)), torch.tensor(np.float32(np.sqrt(var))))
    grads = []
    for r in res:
        agent = oracle_nn_ac.AcAgent(OBS, ACT, 1.0, -1.0)
        agent.load_flat(r["p0"])
        opt = torch.optim.Adam(agent.parameters(), lr=0.0, eps=1e-5)
        tb = {k: torch.from_numpy(r[k]) for k in ("obs", "actions", "logprobs", "advantages", "returns", "values")}
        out = oracle_nn_ac.ac_minibatch_step(agent, opt, tb["obs"], tb["actions"], tb["logprobs"], tb["advantages"], tb["returns"], tb["values"],
                                             torch.from_numpy(r["idx"].astype(np.int64)), adv_moments=moments)
        grads.append(out["grad"])
    g_ref = (grads[0] + grads[1]) * np.float32(0.5)
    g = res[0]["grad"]
    rel = np.abs(g - g_ref).max() / np.abs(g_ref).max()
    assert rel < 2e-4, rel
    # local moments instead of the cross-rank ones must be told apart (the ranks' advantages differ in scale and offset)
    agent = oracle_nn_ac.AcAgent(OBS, ACT, 1.0, -1.0)
    agent.load_flat(res[0]["p0"])
    opt = torch.optim.Adam(agent.parameters(), lr=0.0, eps=1e-5)
    tb = {k: torch.from_numpy(res[0][k]) for k in ("obs", "actions", "logprobs", "advantages", "returns", "values")}
    g_local = oracle_nn_ac.ac_minibatch_step(agent, opt, tb["obs"], tb["actions"], tb["logprobs"], tb["advantages"], tb["returns"], tb["values"],
                                             torch.from_numpy(res[0]["idx"].astype(np.int64)))["grad"]
    assert np.abs(grads[0] - g_local).max() / np.abs(g_local).max() > 1e-2
    # the Adam step on the averaged, clipped gradient
    agent = oracle_nn_ac.AcAgent(OBS, ACT, 1.0, -1.0)
    agent.load_flat(res[0]["p0"])
    opt = torch.optim.Adam(agent.parameters(), lr=2.5e-4, eps=1e-5)
    off = 0
    for p in agent.ordered():
        n = p.numel()
        p.grad = torch.from_numpy(g_ref[off:off + n].copy()).reshape(p.shape)
        off += n
    torch.nn.utils.clip_grad_norm_(agent.parameters(), 0.5)
    opt.step()
    assert np.abs(res[0]["params"] - agent.flat()).max() < 2e-5
