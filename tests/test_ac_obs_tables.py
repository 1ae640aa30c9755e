"""The reference's fixed observation mean / std tables of the ac_ppo agent (src/ac_ppo_continuous_action.cpp:496-497 Humanoid-v4,
:521-522 Ant-v5, :533-534 Hopper-v5) are the only data in the reference that its author measured with the real MuJoCo 3.2.0.
They are carried as data (ppo_cpp_b200/data/ac_obs_stats.json, host/ac_obs_stats.h; tools/extract_ac_obs_stats.py) and pin the
oracle where they can (VERDICT r1 "Next round" 2a):
  * every zero-std entry is a model constant the oracle must reproduce exactly (Humanoid body masses in cinert[:, 9], the world
    body's cvel), every std = 1.0 placeholder with mean 0 is an observation that is identically zero in the author's data
    (world rows, root-dof actuator forces, Ant's jointless *_leg bodies' cfrc_ext);
  * loose sanity bands under a trained policy: the mean height of the walking states sits within the author's mean +- 3 std.
"""
import json
import os

import numpy as np
import pytest

from oracle import pbo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TABLES = json.load(open(os.path.join(ROOT, "ppo_cpp_b200", "data", "ac_obs_stats.json")))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def _rollout_obs(env, n=64, steps=6):
    """Raw (unnormalised) observations of the oracle from the harvested walking states under their recorded actions."""
    d = np.load(os.path.join(GOLDEN, f"walking_{env}.npz"))
    v = pbo.VecEnv(env, n, chain=0)
    v.reset(1)
    for i in range(n):
        s = v.sim(i)
        s.qpos[:] = d["qpos"][i]
        s.qvel[:] = d["qvel"][i]
        s.qacc_warmstart[:] = d["warm"][i]
    out = []
    rng = np.random.default_rng(0)
    for t in range(steps):
        act = d["act"][rng.permutation(len(d["act"]))[:n]]
        o, _, te, tr, _ = v.step(act)
        keep = (te + tr) == 0          # the step after a done returns the reset observation; fine, but keep the walking ones
        out.append(o[keep].copy())
    return np.concatenate(out)


def test_table_shapes():
    assert [len(TABLES[e]["mean"]) for e in ("HalfCheetah-v5", "Hopper-v5", "Ant-v5", "Humanoid-v4")] == [17, 11, 105, 376]
    assert TABLES["HalfCheetah-v5"]["mean"] == [0.0] * 17 and TABLES["HalfCheetah-v5"]["std"] == [1.0] * 17   # :506-508


def test_humanoid_constant_entries_reproduced():
    mean, std = np.array(TABLES["Humanoid-v4"]["mean"]), np.array(TABLES["Humanoid-v4"]["std"])
    obs = _rollout_obs("Humanoid-v4")
    zero = np.where(std == 0)[0]
    # obs layout (humanoid_v4.h:131-158): qpos[2:] 22 | qvel 23 | cinert 14x10 | cvel 14x6 | qfrc_actuator 23 | cfrc_ext 14x6
    assert list(zero) == [45 + 10 * b + 9 for b in range(1, 14)] + list(range(185, 191))
    assert np.abs(obs[:, zero] - mean[zero]).max() < 5e-5      # 4-digit table values: masses of the 13 bodies, zeros
    assert obs[:, zero].std(axis=0).max() < 1e-6
    placeholder = np.where((std == 1.0) & (mean == 0.0))[0]
    assert list(placeholder) == list(range(45, 55)) + list(range(269, 275)) + list(range(292, 298))
    assert np.all(obs[:, placeholder] == 0.0)                   # world cinert row, root-dof actuator forces, world cfrc_ext


def test_ant_jointless_leg_bodies_carry_no_contact_force():
    mean, std = np.array(TABLES["Ant-v5"]["mean"]), np.array(TABLES["Ant-v5"]["std"])
    placeholder = np.where((std == 1.0) & (mean == 0.0))[0]
    # obs layout (ant_v5.h:182-198): qpos[2:] 13 | qvel 14 | cfrc_ext of bodies 1..13; bodies 2, 5, 8, 11 are the jointless *_leg bodies
    assert list(placeholder) == [27 + 6 * (k - 1) + c for k in (2, 5, 8, 11) for c in range(6)]
    obs = _rollout_obs("Ant-v5")
    assert np.all(obs[:, placeholder] == 0.0)
    contact_blocks = [27 + 6 * (k - 1) + c for k in (4, 7, 10, 13) for c in range(6)]     # the ankle bodies do touch the floor
    assert np.abs(obs[:, contact_blocks]).max() > 0.1


@pytest.mark.parametrize("env,idx", [("Hopper-v5", 0), ("Humanoid-v4", 0)])
def test_height_of_walking_states_within_the_authors_band(env, idx):
    """obs[0] is the torso height in both tasks (hopper_v5.h:172-183, humanoid_v4.h:131-137): author's trained-policy mean 1.2739 +-
    0.1744 (Hopper) and 1.2067 +- 0.0838 (Humanoid); our 50-iteration policy's walking states must sit inside mean +- 3 std."""
    mean, std = TABLES[env]["mean"][idx], TABLES[env]["std"][idx]
    obs = _rollout_obs(env)
    assert abs(obs[:, idx].mean() - mean) < 3 * std, (obs[:, idx].mean(), mean, std)
