"""ac_ppo agent on the GPU vs the torch-fp32 oracle (oracle/oracle_nn_ac.py): LayerNorm encoders + Beta heads forward,
the Beta clipped-loss backward with cross-sample advantage moments, Adam, and the sampling kernel's statistics.
Tolerance: fp32 kernels vs fp32 torch CPU with different summation orders -> 2e-4 on logits/values, rel 2e-4 on gradients
(scaled by the largest entry), 2e-5 on parameters after an Adam step (lr 2.5e-4)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return np.abs(a - b).max() / max(1e-12, np.abs(b).max())


@pytest.fixture(scope="module")
def mods():
    import torch

    from oracle import oracle_nn_ac
    from ppo_cpp_b200 import trainer
    torch.manual_seed(0)
    return torch, oracle_nn_ac, trainer


def _agent(torch, oac, obs, act, rng):
    mean, std = rng.normal(size=obs).astype(np.float32) * 0.3, (0.5 + rng.random(obs)).astype(np.float32)
    agent = oac.AcAgent(obs, act, 1.0, -1.0, mean, std)
    with torch.no_grad():  # move LayerNorm affine parameters off their 1 / 0 init so their gradients are exercised
        for p in agent.parameters():
            p.add_(0.03 * torch.randn_like(p))
    return agent, mean, std


@pytest.mark.parametrize("env,obs,act", [("Hopper-v5", 11, 3), ("HalfCheetah-v5", 17, 6), ("Ant-v5", 105, 8), ("Humanoid-v4", 376, 17)])
def test_forward_logits_and_values_match_torch(mods, env, obs, act):
    torch, oac, trainer = mods
    rng = np.random.default_rng(1)
    agent, mean, std = _agent(torch, oac, obs, act, rng)
    N = 200  # not a multiple of the 64-row GEMM tile nor of the 8-row LayerNorm block
    t = trainer.AcTrainer(env_id=env, num_envs=N, num_steps=4, num_minibatches=4, update_epochs=1, obs_mean=mean, obs_std=std)
    t.set_params(agent.flat())
    x = rng.normal(size=(N, obs)).astype(np.float32)
    t.write("next_obs", x)
    zab, v = t.forward()
    zab, v = zab[: N * 2 * act].reshape(N, 2 * act), v[:N]
    with torch.no_grad():
        xt = torch.from_numpy(x)
        f = agent.actor_mean((xt - agent.mean_) / agent.std_)
        za, zb = agent.dist_alpha(f).numpy(), agent.dist_beta(f).numpy()
        val = agent.get_value(xt).view(-1).numpy()
    assert np.abs(zab[:, :act] - za).max() < 2e-4
    assert np.abs(zab[:, act:] - zb).max() < 2e-4
    assert np.abs(v - val).max() < 2e-4
    t.close()


# sizes: a minibatch of 100 rows (ragged against every tile, below the 128-row tensor-core tile: FFMA products) and one of 300 rows
# (tcgen05 products with a ragged last tile; more rows than persistent blocks in the fused row kernels)
@pytest.mark.parametrize("env,obs,act,N,T,nmb", [("Hopper-v5", 11, 3, 50, 8, 4), ("Humanoid-v4", 376, 17, 50, 8, 4), ("HalfCheetah-v5", 17, 6, 75, 8, 2),
                                                 ("Ant-v5", 105, 8, 75, 8, 2), ("HalfCheetah-v5", 17, 6, 50, 8, 4)])
@pytest.mark.parametrize("ent_coef,clip_vloss,norm_adv", [(0.01, 1, 1), (0.0, 0, 0)])
def test_minibatch_update_matches_torch_autograd(mods, env, obs, act, N, T, nmb, ent_coef, clip_vloss, norm_adv):
    torch, oac, trainer = mods
    rng = np.random.default_rng(2)
    agent, mean, std = _agent(torch, oac, obs, act, rng)
    t = trainer.AcTrainer(env_id=env, num_envs=N, num_steps=T, num_minibatches=nmb, update_epochs=1, ent_coef=ent_coef,
                          clip_vloss=clip_vloss, norm_adv=norm_adv, obs_mean=mean, obs_std=std)
    low, high = -1.0, 1.0
    if env == "Humanoid-v4":
        low, high = -0.4, 0.4
    agent.low, agent.high = low, high
    t.set_params(agent.flat())
    B = N * T
    data = dict(obs=rng.normal(size=(B, obs)), actions=rng.uniform(low, high, size=(B, act)) * 0.98, advantages=rng.normal(size=B),
                returns=rng.normal(size=B))
    data = {k: v.astype(np.float32) for k, v in data.items()}
    # actions at the lower bound hit the 1e-7 clamp of scale_action (:194-203).  (At the upper bound the reference's clamp
    # limit 1 + 1e-7 makes log(1 - a) NaN in libtorch as well; sampled actions never reach it, so it is not a parity case.)
    data["actions"][:8] = low
    with torch.no_grad():
        _, lp, _, val = agent.get_action_and_value(torch.from_numpy(data["obs"]), torch.from_numpy(data["actions"]))
    data["logprobs"] = (lp.numpy() + rng.normal(size=B).astype(np.float32) * 0.08).astype(np.float32)
    data["values"] = (val.view(-1).numpy() + rng.normal(size=B).astype(np.float32) * 0.2).astype(np.float32)
    for k, v in data.items():
        t.write(k, v)
    opt = torch.optim.Adam(agent.parameters(), lr=2.5e-4, eps=1e-5)
    tb = {k: torch.from_numpy(v) for k, v in data.items()}
    idx = rng.permutation(B)[: B // nmb].astype(np.int32)
    for step in range(3):
        ref = oac.ac_minibatch_step(agent, opt, tb["obs"], tb["actions"], tb["logprobs"], tb["advantages"], tb["returns"], tb["values"],
                                    torch.from_numpy(idx.astype(np.int64)), clip_coef=0.1, norm_adv=bool(norm_adv),
                                    clip_vloss=bool(clip_vloss), ent_coef=ent_coef)
        t.update_minibatch(idx, 2.5e-4)
        st = t.read("mb_stats")[:6]
        for got, key in zip(st, ["policy_loss", "value_loss", "entropy", "old_approx_kl", "approx_kl", "clipfrac"]):
            assert abs(got - ref[key]) < 2e-4 * max(1.0, abs(ref[key])), (step, key, got, ref[key])
        g = t.read("grad")
        assert np.isfinite(g).all()
        assert _rel(g, ref["grad"]) < 2e-4, (step, _rel(g, ref["grad"]))
        assert np.abs(t.get_params() - agent.flat()).max() < 2e-5, step
        idx = rng.permutation(B)[: B // nmb].astype(np.int32)
    t.close()


def test_rollout_logprobs_values_and_gae_match_torch(mods):
    torch, oac, trainer = mods
    env, obs, act = "Ant-v5", 105, 8
    rng = np.random.default_rng(3)
    agent, mean, std = _agent(torch, oac, obs, act, rng)
    N, T = 96, 12
    t = trainer.AcTrainer(env_id=env, num_envs=N, num_steps=T, num_minibatches=4, update_epochs=1, obs_mean=mean, obs_std=std)
    t.set_params(agent.flat())
    t.reset(1)
    t.rollout()
    t.gae()
    o = torch.from_numpy(t.read("obs").reshape(T * N, obs))
    a = t.read("actions").reshape(T * N, act)
    assert a.min() >= -1.0 and a.max() <= 1.0
    with torch.no_grad():
        _, lp, _, val = agent.get_action_and_value(o, torch.from_numpy(a))
    # stored log-prob is evaluated at the sampled (0,1) value; re-scaling the env-scale action costs ~1 ulp of the action,
    # which the Beta density amplifies near the borders -> compare away from the borders tightly, everywhere loosely
    s = (a + 1.0) / 2.0
    inner = ((s > 1e-3) & (s < 1 - 1e-3)).all(1)
    d = np.abs(t.read("logprobs") - lp.numpy())
    assert inner.mean() > 0.9
    assert d[inner].max() < 2e-3, d[inner].max()
    assert np.abs(t.read("values") - val.view(-1).numpy()).max() < 2e-4
    # GAE recursion (:759-779) on the device buffers
    rew, done, vals = t.read("rewards").reshape(T, N), t.read("dones").reshape(T, N), t.read("values").reshape(T, N)
    nv, term = t.read("next_value"), t.read("term")
    trunc = t.read("trunc")
    next_done = np.maximum(term, trunc)
    adv = np.zeros((T, N), np.float32)
    last = np.zeros(N, np.float32)
    for k in reversed(range(T)):
        nnt = 1.0 - (next_done if k == T - 1 else done[k + 1])
        nxt = nv if k == T - 1 else vals[k + 1]
        delta = rew[k] + 0.99 * nxt * nnt - vals[k]
        last = delta + 0.99 * 0.95 * nnt * last
        adv[k] = last
    assert np.abs(t.read("advantages").reshape(T, N) - adv).max() < 1e-4
    assert np.abs(t.read("returns").reshape(T, N) - (adv + vals)).max() < 1e-4
    t.close()


def test_beta_sampler_moments(mods):
    """Marsaglia-Tsang gamma pairs -> Beta(alpha, beta): sample mean / variance against the closed forms, per action dim."""
    torch, oac, trainer = mods
    env, obs, act = "Hopper-v5", 11, 3
    N = 32768
    t = trainer.AcTrainer(env_id=env, num_envs=N, num_steps=1, num_minibatches=1, update_epochs=1)
    agent = oac.AcAgent(obs, act, 1.0, -1.0)
    with torch.no_grad():  # constant concentrations: zero the head weights, set the biases
        agent.dist_alpha[0].weight.zero_(); agent.dist_beta[0].weight.zero_()
        agent.dist_alpha[0].bias.copy_(torch.tensor([-2.0, 0.5, 3.0])); agent.dist_beta[0].bias.copy_(torch.tensor([1.0, 0.5, -1.0]))
    t.set_params(agent.flat())
    t.reset(1)
    t.rollout()
    a = (t.read("actions").reshape(N, act) + 1.0) / 2.0
    sp = lambda x: np.log1p(np.exp(x))
    al, be = sp(np.array([-2.0, 0.5, 3.0])) + 1, sp(np.array([1.0, 0.5, -1.0])) + 1
    mean, var = al / (al + be), al * be / ((al + be) ** 2 * (al + be + 1))
    se = np.sqrt(var / N)
    assert (np.abs(a.mean(0) - mean) < 5 * se).all(), (a.mean(0), mean)
    assert (np.abs(a.var(0) - var) < 0.03 * var).all(), (a.var(0), var)
    # evaluation mode: action = Beta mean (:229-231)
    r = t.evaluate(7, 4)
    assert len(r) >= 4 and np.isfinite(r).all()
    t.close()


def test_training_improves_return(mods):
    """A few hundred thousand HalfCheetah steps with the reference's update settings move the mean episodic return up."""
    torch, oac, trainer = mods
    t = trainer.AcTrainer(env_id="HalfCheetah-v5", num_envs=256, num_steps=128, num_minibatches=4, update_epochs=4, seed=3)
    t.reset(3)
    first = last = None
    for it in range(24):
        t.iteration(2.5e-4)
        s = t.stats()
        if s["episodes"]:
            m = s["episodic_return_sum"] / s["episodes"]
            first = m if first is None else first
            last = m
        assert np.isfinite(s["policy_loss"]) and np.isfinite(s["value_loss"])
    assert first is not None and last > first + 20.0, (first, last)
    t.close()


def test_big_tile_gemm_and_wgrad_kernels_keep_parity():
    """The 128 x 128 register-tiled GEMM / weight-gradient kernels (PB_AC_GEMM128, read once per process) against the same
    torch-autograd oracle: the forward and minibatch-update tests above, re-run in a subprocess with the big tiles forced."""
    import os
    import subprocess
    import sys
    env = dict(os.environ, PB_AC_GEMM128="2")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-q", "-m", "gpu", "-k",
                        "forward_logits or minibatch_update or rollout_logprobs"], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]


def test_one_op_kernels_keep_parity():
    """PB_AC_FUSED=0 (read at trainer creation): the one-op-per-kernel path that the fused row kernels of pb_ac_fused.cuh replace."""
    import os
    import subprocess
    import sys
    env = dict(os.environ, PB_AC_FUSED="0")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-q", "-m", "gpu", "-k",
                        "forward_logits or minibatch_update"], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]


def test_preempted_collection_trains_on_the_collected_steps_only():
    """DD-PPO pre-emption (src/ac_ppo_continuous_action.cpp:649-693,703-722,759-779,804-810): a collector that breaks out after taking
    step index k returns k, so k steps are trained on; GAE runs over t = k-1 .. 0 and reads slot k of the values / dones buffers
    - written by the step taken before the break - as its "next" (the loop bootstraps from next_value only at
    t == num_steps - 1, :766); the epochs index the first k * N samples, repeated up to the batch size."""
    from ppo_cpp_b200 import trainer
    N, T, k = 32, 16, 6
    t = trainer.AcTrainer(env_id="Hopper-v5", num_envs=N, num_steps=T, num_minibatches=4, update_epochs=2, seed=2)
    t.reset(2)
    t.iteration(2.5e-4)
    old_v = t.read("values").reshape(T, N).copy()
    old_adv = t.read("advantages").reshape(T, N).copy()
    t.collect(k + 1, preempted=True)
    assert t.train_steps() == k
    t.gae()
    t.sync()
    r, v, d = (t.read(n).reshape(T, N) for n in ("rewards", "values", "dones"))
    assert not np.array_equal(v[k], old_v[k]) and np.array_equal(v[k + 1], old_v[k + 1])   # slot k is fresh, k+1 untouched
    adv = np.zeros((k, N), np.float32)
    last = np.zeros(N, np.float32)
    for s in reversed(range(k)):              # :764-776 with num_train_steps_per_env = k
        nnt, nv = 1.0 - d[s + 1], v[s + 1]
        delta = r[s] + 0.99 * nv * nnt - v[s]
        last = delta + 0.99 * 0.95 * nnt * last
        adv[s] = last
    got = t.read("advantages").reshape(T, N)
    np.testing.assert_allclose(got[:k], adv, rtol=1e-5, atol=1e-5)
    assert np.array_equal(got[k:], old_adv[k:])                                  # steps beyond k keep last iteration's values
    p0 = t.get_params()
    t.update(2.5e-4)
    st = t.stats()
    assert np.isfinite(st["policy_loss"]) and np.isfinite(st["value_loss"]) and not np.array_equal(t.get_params(), p0)
    t.rollout()                               # the next full collection resets the step count
    assert t.train_steps() == T
    t.close()
