"""Model compiler known answers (SURVEY.md §8c ii): the only constants the reference itself pins."""
import os

import numpy as np
import pytest

from oracle import pbo
from ppo_cpp_b200 import mjcf_compile as mc

ASSETS = "/root/reference/libs/gymcpp/mujoco/assets"
ENVS = ["HalfCheetah-v5", "Hopper-v5", "Ant-v5", "Humanoid-v4"]
# shape asserts of the task headers: half_cheetah_v5.h:43, hopper_v5.h:50, ant_v5.h:52, humanoid_v4.h:39
DIMS = {"HalfCheetah-v5": (9, 9, 6), "Hopper-v5": (6, 6, 3), "Ant-v5": (15, 14, 8), "Humanoid-v4": (24, 23, 17)}
# zero-std entries of the author's measured obs statistics = body masses (src/ac_ppo_continuous_action.cpp:496-497)
HUMANOID_MASSES = [8.9075, 2.2619, 6.6162, 4.7518, 2.7557, 1.7671, 4.7518, 2.7557, 1.7671, 1.6611, 1.2295, 1.6611,
                   1.2295]


@pytest.mark.parametrize("env", ENVS)
def test_dims_match_reference_asserts(env):
    s = pbo.Sim(env)
    assert (s.nq, s.nv, s.nu) == DIMS[env]


def test_humanoid_body_masses_match_reference_constants():
    s = pbo.Sim("Humanoid-v4")
    s.forward()
    masses = s.cinert.reshape(-1, 10)[1:, 9]
    np.testing.assert_allclose(masses, HUMANOID_MASSES, atol=5e-5)
    assert np.all(s.cinert.reshape(-1, 10)[0] == 0)  # world row is all zeros in the reference data


def test_halfcheetah_settotalmass():
    s = pbo.Sim("HalfCheetah-v5")
    s.forward()
    assert abs(s.cinert.reshape(-1, 10)[1:, 9].sum() - 14.0) < 1e-12


@pytest.mark.skipif(not os.path.isdir(ASSETS), reason="reference assets not present (GPU box)")
@pytest.mark.parametrize("env,xml", list(zip(ENVS, ["half_cheetah", "hopper", "ant", "humanoid"])))
def test_generated_tables_are_current_and_mass_matrix_agrees(env, xml):
    """Committed model_tables.h == fresh compile; oracle CRB mass matrix == independent Jacobian-sum one."""
    md = mc.compile_model(mc.parse_mjcf(f"{ASSETS}/{xml}.xml"))
    s = pbo.Sim(env)
    rng = np.random.default_rng(0)
    s.qpos[:] = md["qpos0"] + rng.uniform(-0.3, 0.3, md["nq"])
    s.forward()
    M_ref, _ = mc.mass_matrix_np(md, s.qpos.copy())
    np.testing.assert_allclose(s.mass_matrix(), M_ref, atol=1e-12)
    assert s.npair == md["npair"]


def test_collision_pair_counts():
    # floor pairs: HC 8 capsules, Hopper 4, Ant 13, Humanoid 17
    assert pbo.Sim("HalfCheetah-v5").npair == 8
    assert pbo.Sim("Hopper-v5").npair == 4 + 3
    assert pbo.Sim("Ant-v5").npair == 13
    assert pbo.Sim("Humanoid-v4").npair == 17 + 109
