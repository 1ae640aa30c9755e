"""Edge cases and full-size properties of the CUDA path (through the C ABI), against the oracle where it finishes in
seconds and through size-independent properties at BASELINE.json's sizes:
  * ragged sizes: minibatches that are not a multiple of any tile height (tcgen05 row tiles of 128, forward/backward tiles
    of 64, weight-gradient splits with empty and partial chunks), env counts that are not a multiple of the CTA's 8 warps,
    a single env (the reference's default `--num_envs 1`, src/ppo_continuous_action.cpp:57);
  * 8192 envs x 32 steps (the bench workload): the minibatch gradient does not depend on the order of the rows inside
    the minibatch (fp32 reassociation only: rel 2e-5), an optimiser step with lr = 0 leaves the parameters bit-identical,
    and two updates from the same state are bit-reproducible (fixed summation order of the split slabs).
Tolerances as in test_gpu_ppo_parity.py: rel 1e-4 on gradients vs torch autograd, 2e-5 abs on parameters."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return np.abs(a - b).max() / max(1e-12, np.abs(b).max())


def _fill(t, rng, B, obs, act, agent=None, torch=None):
    data = dict(obs=rng.normal(size=(B, obs)), actions=rng.normal(size=(B, act)) * 0.5, logprobs=rng.normal(size=B) * 0.3 - act,
                advantages=rng.normal(size=B), returns=rng.normal(size=B), values=rng.normal(size=B) * 0.5)
    data = {k: v.astype(np.float32) for k, v in data.items()}
    if agent is not None:
        with torch.no_grad():
            _, lp, _, val = agent.get_action_and_value(torch.from_numpy(data["obs"]), torch.from_numpy(data["actions"]))
        data["logprobs"] = (lp.numpy() + rng.normal(size=B).astype(np.float32) * 0.15).astype(np.float32)
        data["values"] = (val.view(-1).numpy() + rng.normal(size=B).astype(np.float32) * 0.2).astype(np.float32)
    for k, v in data.items():
        t.write(k, v)
    return data


@pytest.mark.parametrize("env,obs,act,N,T,nmb", [
    ("Humanoid-v4", 376, 17, 50, 7, 2),     # mb = 175: one full 128-row tensor-core tile + 47 rows, 2 full + 1 partial 64-row tile
    ("Humanoid-v4", 376, 17, 3, 5, 1),      # mb = 15: fewer rows than weight-gradient splits (empty chunks), one partial tile
    ("Hopper-v5", 11, 3, 33, 9, 3),         # mb = 99
    ("Ant-v5", 105, 8, 1, 40, 2),           # a single env, mb = 20
])
def test_ragged_minibatch_update_matches_torch_autograd(env, obs, act, N, T, nmb):
    import torch

    from oracle import oracle_nn as onn
    from ppo_cpp_b200 import trainer
    torch.manual_seed(1)
    t = trainer.Trainer(env_id=env, num_envs=N, num_steps=T, num_minibatches=nmb, update_epochs=1)
    agent = onn.Agent(obs, act)
    with torch.no_grad():
        agent.actor_logstd.normal_(0, 0.3)
        for p in agent.parameters():
            p.add_(0.05 * torch.randn_like(p))
    t.set_params(agent.flat())
    rng = np.random.default_rng(N * 1000 + T)
    B = N * T
    data = _fill(t, rng, B, obs, act, agent, torch)
    tb = {k: torch.from_numpy(v) for k, v in data.items()}
    opt = torch.optim.Adam(agent.parameters(), lr=3e-4, eps=1e-5)
    for step in range(2):
        idx = rng.permutation(B)[: B // nmb].astype(np.int32)
        ref = onn.ppo_minibatch_step(agent, opt, tb["obs"], tb["actions"], tb["logprobs"], tb["advantages"], tb["returns"],
                                     tb["values"], torch.from_numpy(idx.astype(np.int64)))
        t.update_minibatch(idx, 3e-4)
        assert _rel(t.read("grad"), ref["grad"]) < 1e-4, step
        assert np.abs(t.get_params() - agent.flat()).max() < 2e-5, step
    t.close()


@pytest.mark.parametrize("env,n", [("Humanoid-v4", 13), ("Hopper-v5", 1), ("Ant-v5", 9)])
def test_env_counts_that_do_not_fill_a_cta(env, n):
    from oracle import pbo
    from ppo_cpp_b200 import capi
    gv, ov = capi.ParVectorEnv(env, n, chain=capi.PB_CHAIN_STATS_ONLY), pbo.VecEnv(env, n, chain=0)
    g0, o0 = gv.reset(5).copy(), ov.reset(5).copy()
    assert np.abs(g0 - o0).max() < 1e-4
    rng = np.random.default_rng(n)
    act = rng.uniform(-0.4, 0.4, size=(n, gv.act_dim)).astype(np.float32)
    g = [x.copy() for x in gv.step(act)[:4]]
    o = [x.copy() for x in ov.step(act)[:4]]
    rel = np.abs(g[0] - o[0]) / np.maximum(1.0, np.abs(o[0]))
    assert np.median(rel.max(axis=1)) < 1e-3 and rel.max() < 1e-2
    assert np.abs(g[1] - o[1]).max() < 1e-2 and np.array_equal(g[2], o[2]) and np.array_equal(g[3], o[3])
    assert gv.counters()[1] == 0
    gv.close()


def test_single_env_iterations_like_the_reference_default():
    from ppo_cpp_b200 import trainer
    t = trainer.Trainer(env_id="HalfCheetah-v5", num_envs=1, num_steps=64, num_minibatches=4, update_epochs=2)
    t.reset(1)
    for _ in range(2):
        t.iteration(3e-4)
    st = t.stats()
    assert np.isfinite([st["value_loss"], st["policy_loss"], st["entropy"], st["approx_kl"]]).all()
    assert np.isfinite(t.get_params()).all()
    t.close()


def test_full_size_minibatch_properties():
    """The bench workload's sizes: 8192 envs x 32 steps, minibatch 8192 rows of 376 observations."""
    from ppo_cpp_b200 import trainer
    N, T, nmb, obs, act = 8192, 32, 32, 376, 17
    t = trainer.Trainer(env_id="Humanoid-v4", num_envs=N, num_steps=T, num_minibatches=nmb, update_epochs=1, seed=2)
    rng = np.random.default_rng(0)
    B, mb = N * T, N * T // nmb
    _fill(t, rng, B, obs, act)
    p0 = t.get_params()
    idx = rng.permutation(B)[:mb].astype(np.int32)
    t.update_minibatch(idx, 0.0)
    g1 = t.read("grad")
    assert np.array_equal(t.get_params(), p0)                       # lr = 0: bit-identical parameters
    t.update_minibatch(idx, 0.0)
    assert np.array_equal(t.read("grad"), g1)                       # bit-reproducible (fixed slab summation order)
    t.update_minibatch(rng.permutation(idx).astype(np.int32), 0.0)
    g2 = t.read("grad")
    assert np.isfinite(g1).all() and np.abs(g1).max() > 0
    assert _rel(g2, g1) < 2e-5                                      # row order inside the minibatch: reassociation only
    t.close()


def test_clipfrac_is_the_mean_over_the_iterations_minibatches():
    """losses/clipfrac (src/ppo_continuous_action.cpp:505,581) is the mean of the clipfracs of all update_epochs x num_minibatches
    minibatches of the iteration, not the last minibatch's.  Trainer A runs the whole update; trainer B replays the
    same minibatches one at a time (same seed -> same parameters, rollout and host permutation stream)."""
    from ppo_cpp_b200 import capi, trainer
    kw = dict(env_id="Hopper-v5", num_envs=64, num_steps=16, num_minibatches=4, update_epochs=3, seed=5)
    a, b = trainer.Trainer(**kw), trainer.Trainer(**kw)
    for t in (a, b):
        t.reset(5)
        t.iteration(3e-3)          # one real update first, so that ratios leave the clip range in the next one
        t.rollout()
        t.gae()
    a.update(3e-3)
    sa = a.stats()
    B, mb = 64 * 16, 64 * 16 // 4
    rng = C.c_void_p(capi.lib().pb_rng_create(5))
    for _ in range(3):             # the first update consumed three permutations
        trainer.randperm(rng, B)
    fracs = []
    for e in range(3):
        perm = trainer.randperm(rng, B)
        for k in range(4):
            b.update_minibatch(perm[k * mb:(k + 1) * mb].astype(np.int32), 3e-3)
            fracs.append(b.stats()["clipfrac"])
    capi.lib().pb_rng_destroy(rng)
    assert np.array_equal(a.get_params(), b.get_params())          # the replay really was the same update
    assert max(fracs) > 0 and len(set(np.round(fracs, 6))) > 1     # the minibatches differ, so mean != last
    assert abs(sa["clipfrac"] - np.mean(fracs)) < 1e-6
    assert abs(sa["clipfrac"] - fracs[-1]) > 1e-4
    a.close(); b.close()


def test_trainers_on_two_devices_in_one_process():
    """cudaFuncSetAttribute state is per device: a second trainer on device 1 must be able to launch the kernels
    that need > 48 KB of dynamic shared memory."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from ppo_cpp_b200 import trainer
    for dev in (0, 1):
        t = trainer.Trainer(env_id="Humanoid-v4", num_envs=64, num_steps=8, num_minibatches=4, update_epochs=1, device=dev)
        t.reset(1)
        t.iteration(3e-4)
        assert np.isfinite(t.stats()["value_loss"])
        t.close()
