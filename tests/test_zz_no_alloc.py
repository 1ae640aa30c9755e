"""pb_obs_moments_noalloc (include/ppo_b200.h): the allocation-free variant of the observation-moment reduction, checked
against the existing entry point pb_ac_trainer_obs_sums through the C ABI, on two calls in a row (the second with another
shape, so a stale partial of the first call left in the shared static scratch would show up).  No counterpart in the
reference.  Collected last on purpose.

Tolerance: both paths accumulate in fp64 and differ only in summation order -> rel 1e-10 of the column's sum of |x|.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _reference_sums(trainer, capi, env_id, num_envs, num_steps, seed):
    """Existing entry point on a real rollout's observations; returns (obs rows [T*N, O] float32, sums [2*O] float64)."""
    t = trainer.AcTrainer(env_id=env_id, num_envs=num_envs, num_steps=num_steps, num_minibatches=4, update_epochs=1)
    t.reset(seed)
    t.rollout()
    O = capi.lib().pb_ac_trainer_obs_dim(t._h)
    sums, rows = np.zeros(2 * O), C.c_longlong()
    capi._check(capi.lib().pb_ac_trainer_obs_sums(t._h, sums.ctypes.data_as(capi._dp), C.byref(rows)))
    obs = t.read("obs").reshape(-1, O).copy()
    assert rows.value == obs.shape[0] == num_envs * num_steps
    t.close()
    return obs, sums


def _noalloc_sums(capi, torch, obs):
    x = torch.from_numpy(obs).cuda().contiguous()
    out = torch.full((2 * obs.shape[1],), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    st = torch.cuda.current_stream().cuda_stream
    capi._check(capi.lib().pb_obs_moments_noalloc(0, C.c_void_p(st), C.c_void_p(x.data_ptr()), x.shape[0], x.shape[1],
                                                  C.c_void_p(out.data_ptr())))
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_noalloc_matches_existing_entry_point_twice():
    import torch

    from ppo_cpp_b200 import capi, trainer
    # call 1: Humanoid-v4 (376 columns, 64 x 8 rows); call 2 right after it: Hopper-v5 (11 columns, 48 x 5 rows - fewer rows than
    # reduction slices, so some slices are empty and must overwrite the first call's partials with zeros)
    for env_id, n, T, seed in (("Humanoid-v4", 64, 8, 1), ("Hopper-v5", 48, 5, 2)):
        obs, ref = _reference_sums(trainer, capi, env_id, n, T, seed)
        got = _noalloc_sums(capi, torch, obs)
        assert np.all(np.isfinite(got))
        scale = np.repeat(np.abs(obs.astype(np.float64)).sum(axis=0), 2)
        scale[1::2] = (obs.astype(np.float64) ** 2).sum(axis=0)
        assert np.all(np.abs(got - ref) <= 1e-10 * np.maximum(1.0, scale)), (env_id, np.abs(got - ref).max())
        # and against numpy directly (the existing entry point is itself checked this way)
        np.testing.assert_allclose(got[0::2], obs.astype(np.float64).sum(axis=0), rtol=0, atol=1e-10 * np.maximum(1.0, scale[0::2]).max())


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch

    from ppo_cpp_b200 import capi
    L = capi.lib()
    mx = L.pb_obs_moments_noalloc_max_obs()
    assert mx == 512
    x = torch.zeros(4, mx + 1, device="cuda")
    out = torch.zeros(2 * (mx + 1), dtype=torch.float64, device="cuda")
    rc = L.pb_obs_moments_noalloc(0, None, C.c_void_p(x.data_ptr()), 4, mx + 1, C.c_void_p(out.data_ptr()))
    assert rc != 0 and b"exceeds the static scratch limit" in L.pb_last_error()
    # the limit itself is accepted
    x = torch.ones(4, mx, device="cuda")
    out = torch.zeros(2 * mx, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    assert L.pb_obs_moments_noalloc(0, None, C.c_void_p(x.data_ptr()), 4, mx, C.c_void_p(out.data_ptr())) == 0
    torch.cuda.synchronize()
    assert torch.all(out == 4.0)
