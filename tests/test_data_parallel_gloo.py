"""Data-parallel semantics on the CPU (gloo, world_size 2): the engine shards envs across ranks and averages the flat
gradient once per minibatch (src/ac_ppo_continuous_action.cpp:878-885, src/distributed.cpp:117-133).  With equal shard sizes
the average of the per-rank gradients is the gradient of the union minibatch, rank 0's parameters are what every rank starts
from after the broadcast (:551-553), and each rank draws its own permutation stream (seed*1500 + rank, :618).  The oracle's
torch step stands in for the kernels; the NCCL path itself is exercised by `bench.py --gpus N`."""
import os
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, initfile, out):
    from oracle import oracle_nn
    from ppo_cpp_b200 import trainer
    dist.init_process_group("gloo", init_method=f"file://{initfile}", rank=rank, world_size=world)
    torch.manual_seed(100 + rank)  # different init per rank on purpose
    obs, act, B = 11, 3, 128
    agent = oracle_nn.Agent(obs, act)
    flat = torch.from_numpy(agent.flat())
    dist.broadcast(flat, src=0)  # pb_trainer_broadcast_params
    agent.load_flat(flat.numpy())
    rng = np.random.default_rng(0)  # same full batch on every rank; rank r owns rows [r*B/world, (r+1)*B/world)
    full = dict(obs=rng.normal(size=(B, obs)), actions=rng.normal(size=(B, act)) * 0.5, logprobs=rng.normal(size=B) * 0.1 - act,
                adv=rng.normal(size=B), ret=rng.normal(size=B), val=rng.normal(size=B) * 0.5)
    full = {k: torch.from_numpy(v.astype(np.float32)) for k, v in full.items()}
    lo, hi = rank * B // world, (rank + 1) * B // world
    opt = torch.optim.SGD(agent.parameters(), lr=0.0)
    idx = torch.arange(lo, hi)
    ref = oracle_nn.ppo_minibatch_step(agent, opt, full["obs"], full["actions"], full["logprobs"], full["adv"], full["ret"], full["val"],
                                       idx, norm_adv=False, max_grad_norm=1e9)
    g = torch.from_numpy(ref["grad"].copy())
    dist.all_reduce(g, op=dist.ReduceOp.SUM)
    g /= world  # ncclAvg
    perm = trainer.randperm(7 * 1500 + rank, 64)
    if rank == 0:
        np.savez(out, grad=g.numpy(), params=agent.flat(), perm=perm)
    else:
        np.savez(out + ".r1", perm=perm, params=agent.flat())
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_gradient_average_equals_union_minibatch():
    from oracle import oracle_nn
    with tempfile.TemporaryDirectory() as td:
        initfile, out = os.path.join(td, "init"), os.path.join(td, "out.npz")
        mp.spawn(_worker, args=(2, initfile, out), nprocs=2, join=True)
        r0, r1 = np.load(out), np.load(out + ".r1.npz")
    assert np.array_equal(r0["params"], r1["params"])          # broadcast made the replicas identical
    assert not np.array_equal(r0["perm"], r1["perm"])          # per-rank permutation streams
    obs, act, B = 11, 3, 128
    agent = oracle_nn.Agent(obs, act)
    agent.load_flat(r0["params"])
    rng = np.random.default_rng(0)
    full = dict(obs=rng.normal(size=(B, obs)), actions=rng.normal(size=(B, act)) * 0.5, logprobs=rng.normal(size=B) * 0.1 - act,
                adv=rng.normal(size=B), ret=rng.normal(size=B), val=rng.normal(size=B) * 0.5)
    full = {k: torch.from_numpy(v.astype(np.float32)) for k, v in full.items()}
    opt = torch.optim.SGD(agent.parameters(), lr=0.0)
    ref = oracle_nn.ppo_minibatch_step(agent, opt, full["obs"], full["actions"], full["logprobs"], full["adv"], full["ret"], full["val"],
                                       torch.arange(B), norm_adv=False, max_grad_norm=1e9)
    assert np.abs(r0["grad"] - ref["grad"]).max() / np.abs(ref["grad"]).max() < 1e-5
