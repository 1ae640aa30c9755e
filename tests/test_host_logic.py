"""CPU-side checks: bit-exact minibatch permutations, the std::mt19937/libstdc++ streams of the oracle, the wrapper
arithmetic, and that the C-ABI library loads and exports every symbol include/ppo_b200.h declares."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from oracle import pbo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _lib_or_skip():
    from ppo_cpp_b200 import capi
    if not os.path.exists(capi.LIB_PATH):
        pytest.skip("libppo_b200.so not built (run __graft_entry__.build())")
    return capi


def test_capi_exports_every_declared_symbol():
    capi = _lib_or_skip()
    hdr = open(os.path.join(ROOT, "include", "ppo_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(pb_[a-z0-9_]+)\s*\(", hdr))
    L = capi.lib()
    missing = [s for s in sorted(declared) if not hasattr(L, s)]
    assert not missing, missing
    assert declared <= set(capi.SIGNATURES), sorted(declared - set(capi.SIGNATURES))


@pytest.mark.parametrize("seed,n", [(1, 2), (1, 65536), (7, 1000), (3001, 262144)])
def test_randperm_bit_exact_with_torch(seed, n):
    _lib_or_skip()
    from ppo_cpp_b200 import trainer
    g = torch.Generator()
    g.manual_seed(seed)
    assert np.array_equal(torch.randperm(n, generator=g).numpy(), trainer.randperm(seed, n))


def test_randperm_stream_continues_across_epochs():
    capi = _lib_or_skip()
    from ppo_cpp_b200 import trainer
    g = torch.Generator()
    g.manual_seed(1 * 1500 + 0)  # ac_ppo train_generator seed (src/ac_ppo_continuous_action.cpp:618)
    r = C.c_void_p(capi.lib().pb_rng_create(1500))
    for _ in range(4):
        assert np.array_equal(torch.randperm(4096, generator=g).numpy(), trainer.randperm(r, 4096))
    capi.lib().pb_rng_destroy(r)


def test_oracle_mt19937_matches_numpy_legacy_stream():
    L = pbo.lib()
    class Rng(C.Structure):
        _fields_ = [("mt", C.c_uint32 * 624), ("idx", C.c_int), ("saved_ok", C.c_int), ("saved", C.c_double)]
    r = Rng()
    L.pbo_rng_seed(C.byref(r), 12345)
    L.pbo_rng_u32.restype = C.c_uint32
    got = [L.pbo_rng_u32(C.byref(r)) for _ in range(1500)]
    ref = np.random.RandomState(12345).randint(0, 2 ** 32, size=1500, dtype=np.uint64)
    assert got == [int(x) for x in ref]


def test_reset_noise_ranges_and_seed_rule():
    for env, scale in [("Hopper-v5", 5e-3), ("Humanoid-v4", 1e-2)]:
        v = pbo.VecEnv(env, 8, chain=0)
        a = v.reset(5).copy()
        b = v.reset(5).copy()
        assert np.array_equal(a, b)
        c = v.reset(6).copy()
        assert np.array_equal(c[:-1], a[1:])  # env i uses seed+i (gym.h:313)
        s = v.sim(0)
        # qvel noise is U(-scale, scale) around zero
        assert np.abs(a[:, s.nq - (1 if env == "Hopper-v5" else 2):][:, : s.nv]).max() <= scale + 1e-9


def test_wrapper_chain_matches_line_by_line_numpy():
    """NormalizeObservation / NormalizeReward (float32 Welford, count as float32) against a numpy restatement."""
    v = pbo.VecEnv("Hopper-v5", 1, chain=1)
    raw = pbo.VecEnv("Hopper-v5", 1, chain=0)
    f = np.float32
    mean, var, count = np.zeros(11, f), np.ones(11, f), f(1e-4)
    rmean, rvar, racc, rcount = f(0), f(1), f(0), f(1e-8)

    def upd(x, mean, var, count):
        delta = x - mean
        tot = f(count + f(1))
        new_mean = mean + delta * f(1) / tot
        M2 = var * count + f(0) + (delta * delta) * count * f(1) / tot
        return new_mean, (M2 / tot).astype(f) if isinstance(M2, np.ndarray) else f(M2 / tot), tot

    o, r0 = v.reset(3).copy(), raw.reset(3).copy()
    mean, var, count = upd(r0[0], mean, var, count)
    exp = np.clip((r0[0] - mean) / np.sqrt(var + f(1e-4)), -10, 10)
    np.testing.assert_allclose(o[0], exp, rtol=0, atol=1e-6)
    rng = np.random.default_rng(0)
    for t in range(30):
        a = rng.normal(size=(1, 3)).astype(f)
        o, r, te, tr, _ = v.step(a)
        ro, rr, rte, rtr, _ = raw.step(a)
        if v.envs_autoreset if hasattr(v, "envs_autoreset") else False:
            break
        mean, var, count = upd(ro[0], mean, var, count)
        exp = np.clip((ro[0] - mean) / np.sqrt(var + f(1e-4)), -10, 10)
        np.testing.assert_allclose(o[0], exp, rtol=0, atol=2e-6)
        if rr[0] == 0 and rte[0] == 0 and rtr[0] == 0 and t > 0 and prev_done:
            assert r[0] == 0  # autoreset step: reward bypasses the wrappers
        else:
            racc = f(f(racc * f(0.99)) * f(1 - rte[0]) + rr[0])
            rmean, rvar, rcount = upd(racc, rmean, rvar, rcount)
            assert abs(r[0] - np.clip(rr[0] / np.sqrt(rvar + f(1e-8)), -10, 10)) < 2e-6
        prev_done = bool(rte[0] or rtr[0])


def test_dd_ppo_index_rule_bit_exact_with_torch():
    """src/ac_ppo_continuous_action.cpp:804-810: b_inds = randperm(num_collected, generator); under pre-emption
    b_inds.repeat(ceil(batch / num_collected))[:batch].  Successive epochs continue the same generator."""
    import ctypes as C

    import torch

    from ppo_cpp_b200 import capi, trainer
    seed = 3 * 1500 + 1          # the per-rank generator seed of :618
    g = torch.Generator().manual_seed(seed)
    rng = C.c_void_p(capi.lib().pb_rng_create(seed))
    for collected, batch in ((1024, 1024), (640, 1024), (96, 1024), (1023, 1024), (1, 7)):
        for _epoch in range(2):
            b = torch.randperm(collected, generator=g, dtype=torch.long)
            rep = (batch + collected - 1) // collected
            ref = b.repeat(rep)[:batch].numpy()
            got = trainer.dd_ppo_batch_indices(rng, collected, batch)
            assert np.array_equal(got, ref), (collected, batch)
    capi.lib().pb_rng_destroy(rng)
    with pytest.raises(capi.PbError):
        trainer.dd_ppo_batch_indices(C.c_void_p(capi.lib().pb_rng_create(1)), 2048, 1024)


def test_dd_ppo_preempt_rule():
    """:680-688 — (num_done / num_envs > threshold) and step > lround(min_perc * num_steps), in float arithmetic."""
    from ppo_cpp_b200 import capi
    f = capi.lib().pb_dd_ppo_should_preempt
    for num_done, num_envs, thr, step, perc, T in [(7, 10, 0.6, 40, 0.25, 128), (6, 10, 0.6, 40, 0.25, 128), (7, 10, 0.6, 32, 0.25, 128),
                                                   (7, 10, 0.6, 33, 0.25, 128), (61, 100, 0.6, 4, 0.03, 100), (61, 100, 0.6, 3, 0.03, 100),
                                                   (5, 8, 0.625, 100, 0.25, 128), (0, 8, 0.0, 100, 0.25, 128), (1, 8, 0.0, 100, 0.25, 128)]:
        ref = (np.float32(num_done) / np.float32(num_envs) > np.float32(thr)) and step > int(np.floor(np.float32(perc) * np.float32(T) + 0.5))
        assert bool(f(num_done, num_envs, thr, step, perc, T)) == bool(ref), (num_done, num_envs, thr, step, perc, T)
