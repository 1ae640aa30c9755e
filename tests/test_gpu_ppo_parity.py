"""Agent forward / sampling / GAE / clipped-loss backward / Adam on the GPU vs the torch-fp32 oracle (oracle_nn.py).
Tolerance: fp32 kernels vs fp32 torch CPU, different summation orders -> rel 1e-4 on gradients (|g| scaled by the
largest entry), 2e-5 on parameters after one Adam step, 1e-4 on log-probs/values."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return np.abs(a - b).max() / max(1e-12, np.abs(b).max())


@pytest.fixture(scope="module")
def mods():
    import torch

    from oracle import oracle_nn
    from ppo_cpp_b200 import trainer
    torch.manual_seed(0)
    return torch, oracle_nn, trainer


@pytest.mark.parametrize("env,obs,act", [("Hopper-v5", 11, 3), ("Humanoid-v4", 376, 17)])
@pytest.mark.parametrize("ent_coef,clip_vloss", [(0.0, 1), (0.01, 0)])
def test_minibatch_update_matches_torch_autograd(mods, env, obs, act, ent_coef, clip_vloss):
    torch, onn, trainer = mods
    N, T, nmb = 64, 8, 4
    t = trainer.Trainer(env_id=env, num_envs=N, num_steps=T, num_minibatches=nmb, update_epochs=1, ent_coef=ent_coef,
                        clip_vloss=clip_vloss)
    agent = onn.Agent(obs, act)
    with torch.no_grad():
        agent.actor_logstd.normal_(0, 0.3)
        for p in agent.parameters():
            p.add_(0.05 * torch.randn_like(p))
    t.set_params(agent.flat())
    rng = np.random.default_rng(0)
    B = N * T
    data = dict(obs=rng.normal(size=(B, obs)), actions=rng.normal(size=(B, act)) * 0.5, logprobs=rng.normal(size=B) * 0.3 - act,
                advantages=rng.normal(size=B), returns=rng.normal(size=B), values=rng.normal(size=B) * 0.5)
    data = {k: v.astype(np.float32) for k, v in data.items()}
    # make old log-probs consistent enough that ratios straddle the clip range
    with torch.no_grad():
        _, lp, _, val = agent.get_action_and_value(torch.from_numpy(data["obs"]), torch.from_numpy(data["actions"]))
    data["logprobs"] = (lp.numpy() + rng.normal(size=B).astype(np.float32) * 0.15).astype(np.float32)
    data["values"] = (val.view(-1).numpy() + rng.normal(size=B).astype(np.float32) * 0.2).astype(np.float32)
    for k, v in data.items():
        t.write(k, v)
    opt = torch.optim.Adam(agent.parameters(), lr=3e-4, eps=1e-5)
    tb = {k: torch.from_numpy(v) for k, v in data.items()}
    idx = rng.permutation(B)[: B // nmb].astype(np.int32)
    for step in range(3):  # three consecutive optimiser steps (Adam bias correction, moments)
        ref = onn.ppo_minibatch_step(agent, opt, tb["obs"], tb["actions"], tb["logprobs"], tb["advantages"], tb["returns"],
                                     tb["values"], torch.from_numpy(idx.astype(np.int64)), ent_coef=ent_coef,
                                     clip_vloss=bool(clip_vloss))
        t.update_minibatch(idx, 3e-4)
        g = t.read("grad")
        assert _rel(g, ref["grad"]) < 1e-4, (step, _rel(g, ref["grad"]))
        st = t.read("mb_stats")  # pg, v, entropy, old_kl, kl, clipfrac
        for got, key in zip(st, ["policy_loss", "value_loss", "entropy", "old_approx_kl", "approx_kl", "clipfrac"]):
            assert abs(got - ref[key]) < 1e-4 * max(1.0, abs(ref[key])), (key, got, ref[key])
        assert np.abs(t.get_params() - agent.flat()).max() < 2e-5, step
        idx = rng.permutation(B)[: B // nmb].astype(np.int32)
    t.close()


@pytest.mark.parametrize("env,obs,act", [("Hopper-v5", 11, 3), ("Ant-v5", 105, 8)])
def test_rollout_logprob_value_and_gae_match_torch(mods, env, obs, act):
    torch, onn, trainer = mods
    N, T = 128, 16
    t = trainer.Trainer(env_id=env, num_envs=N, num_steps=T, num_minibatches=4, update_epochs=1)
    agent = onn.Agent(obs, act)
    with torch.no_grad():
        agent.actor_logstd.fill_(-0.5)
    t.set_params(agent.flat())
    t.reset(1)
    t.rollout()
    t.gae()
    o = torch.from_numpy(t.read("obs").reshape(T * N, obs))
    a = torch.from_numpy(t.read("actions").reshape(T * N, act))
    with torch.no_grad():
        _, lp, _, val = agent.get_action_and_value(o, a)
        mean = agent.actor_mean(o)
    assert np.abs(t.read("logprobs") - lp.numpy()).max() < 2e-4
    assert np.abs(t.read("values") - val.view(-1).numpy()).max() < 1e-4
    # sampled noise is standard normal: z = (a - mean) / std
    z = ((a - mean) / np.exp(-0.5)).numpy().ravel()
    assert abs(z.mean()) < 0.02 and abs(z.std() - 1.0) < 0.02 and abs((z ** 3).mean()) < 0.05
    # dones[t] is the done flag of the previous transition (CleanRL convention)
    rew = torch.from_numpy(t.read("rewards").reshape(T, N))
    vals = torch.from_numpy(t.read("values").reshape(T, N))
    dones = torch.from_numpy(t.read("dones").reshape(T, N))
    next_done = torch.from_numpy(np.maximum(t.read("term"), t.read("trunc")))
    adv, ret = onn.gae(rew, vals, dones, torch.from_numpy(t.read("next_value")), next_done, 0.99, 0.95)
    assert np.abs(t.read("advantages").reshape(T, N) - adv.numpy()).max() < 1e-4
    assert np.abs(t.read("returns").reshape(T, N) - ret.numpy()).max() < 1e-4
    with torch.no_grad():
        nv = agent.get_value(torch.from_numpy(t.read("next_obs").reshape(N, obs))).view(-1).numpy()
    assert np.abs(t.read("next_value") - nv).max() < 1e-4
    t.close()


def test_full_iterations_run_and_learn_signal(mods):
    """A few complete iterations on Hopper: losses finite, parameters move, episodes get counted."""
    _, _, trainer = mods
    t = trainer.Trainer(env_id="Hopper-v5", num_envs=256, num_steps=16, num_minibatches=4, update_epochs=4)
    p0 = t.get_params()
    t.reset(1)
    for it in range(3):
        t.iteration(3e-4)
    st = t.stats()
    assert all(np.isfinite(st[k]) for k in ("value_loss", "policy_loss", "entropy", "approx_kl"))
    assert st["episodes"] > 0 and st["overflow_envs"] == 0
    assert np.abs(t.get_params() - p0).max() > 1e-5
    t.close()
