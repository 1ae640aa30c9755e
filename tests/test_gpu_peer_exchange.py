"""Two-GPU data-parallel exchange step (src/ac_ppo_continuous_action.cpp:878-885, src/distributed.cpp:117-133) through the
C ABI: the fused peer-memory all-reduce (csrc/pb_peer.cuh) against (a) the arithmetic mean of the two single-rank
gradients — bit-exact, the kernel sums in rank order and scales by 1/world — and (b) the ncclAllReduce(avg) path
(PB_PEER_AR=0) within fp32 rounding (rel 1e-6 on gradients, 1e-6 abs on parameters after three Adam steps).  Replicas must
stay bit-identical.  Needs two devices on one node; skipped otherwise."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import os, sys, time
import numpy as np
sys.path.insert(0, sys.argv[1])
rank, world, td = int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
from ppo_cpp_b200 import trainer
env, obs, act, N, T, nmb = "Humanoid-v4", 376, 17, 64, 8, 4
B = N * T
def make(world_size):
    t = trainer.Trainer(env_id=env, num_envs=N, num_steps=T, num_minibatches=nmb, update_epochs=1, device=rank, rank=rank,
                        world_size=world_size, seed=3)
    rng = np.random.default_rng(100 + rank)     # every rank owns different rows
    data = dict(obs=rng.normal(size=(B, obs)), actions=rng.normal(size=(B, act)) * 0.5, logprobs=rng.normal(size=B) * 0.1 - 0.5 * act,
                advantages=rng.normal(size=B), returns=rng.normal(size=B), values=rng.normal(size=B) * 0.5)
    for k, v in data.items():
        t.write(k, v.astype(np.float32))
    return t
idx = [np.random.default_rng(7).permutation(B)[: B // nmb].astype(np.int32) for _ in range(3)]
# single-rank gradient of the first step (same seed -> same initial parameters on both ranks)
t1 = make(1)
p0 = t1.get_params()
t1.update_minibatch(idx[0], 3e-4)
g_local = t1.read("grad")
t1.close()
t = make(world)
assert np.array_equal(t.get_params(), p0)
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
t.broadcast_params()
grads, params = [], []
for s in range(3):
    t.update_minibatch(idx[s], 3e-4)
    grads.append(t.read("grad")); params.append(t.get_params())
# a whole update (epochs x minibatches back to back, no host sync in between) exercises the slot reuse
t.update(3e-4)
params.append(t.get_params())
np.savez(os.path.join(td, f"r{rank}.npz"), g_local=g_local, grads=np.stack(grads), params=np.stack(params))
t.close()
"""


def _run(world, peer, **extra):
    with tempfile.TemporaryDirectory() as td:
        script = os.path.join(td, "w.py")
        with open(script, "w") as f:
            f.write(WORKER)
        env = dict(os.environ, PB_PEER_AR="1" if peer else "0", **extra)
        procs = [subprocess.Popen([sys.executable, script, ROOT, str(r), str(world), td], env=env, stderr=subprocess.PIPE, text=True)
                 for r in range(world)]
        errs = []
        for p in procs:
            try:
                _, e = p.communicate(timeout=240)
            except subprocess.TimeoutExpired:
                for q in procs:
                    q.kill()
                raise
            errs.append(e)
        assert all(p.returncode == 0 for p in procs), errs
        return [dict(np.load(os.path.join(td, f"r{r}.npz"))) for r in range(world)], errs


@pytest.mark.timeout(600)
def test_peer_memory_gradient_average_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    peer, errs = _run(2, True)
    assert "fused peer-memory all-reduce" in errs[0], errs[0]       # the peer path really ran (not the NCCL fallback)
    mean = (peer[0]["g_local"] + peer[1]["g_local"]) * np.float32(0.5)
    assert np.array_equal(peer[0]["grads"][0], mean)                 # bit-exact average of the per-rank gradients
    assert np.array_equal(peer[0]["grads"], peer[1]["grads"])
    assert np.array_equal(peer[0]["params"], peer[1]["params"])      # replicas stay bit-identical
    assert np.isfinite(peer[0]["params"]).all()
    assert "{value, seq} words" in errs[0], errs[0]                 # default: the one-kernel LL exchange
    two, errs = _run(2, True, PB_PEER_LL="0")                       # flags + remote reads in a second kernel: same bits
    assert "two kernels" in errs[0], errs[0]
    assert np.array_equal(two[0]["grads"], peer[0]["grads"]) and np.array_equal(two[0]["params"], peer[0]["params"])
    nccl, errs = _run(2, False)
    assert "ncclAllReduce" in errs[0], errs[0]
    scale = np.abs(nccl[0]["grads"]).max()
    assert np.abs(peer[0]["grads"] - nccl[0]["grads"]).max() <= 1e-6 * scale
    assert np.abs(peer[0]["params"][:3] - nccl[0]["params"][:3]).max() <= 1e-6
    assert np.abs(peer[0]["params"][3] - nccl[0]["params"][3]).max() <= 1e-5   # four further Adam steps
