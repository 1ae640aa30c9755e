"""Physical invariants that pin the oracle's physics restatement (parity to MuJoCo itself is unpinned)."""
import numpy as np
import pytest

from oracle.pbo import NO_ARMATURE, NO_CONTACT, NO_GRAVITY, NO_LIMIT, NO_PASSIVE, Sim


def _drift(env, integ, h, T=0.4, vel=3.0):
    s = Sim(env, flags=NO_CONTACT | NO_LIMIT | NO_PASSIVE | NO_ARMATURE, integrator=integ, timestep=h)
    r = np.random.default_rng(2)
    s.qpos[:] = s.qpos + r.uniform(-0.2, 0.2, s.nq)
    s.qvel[:] = r.normal(size=s.nv) * vel
    s.forward()
    e0 = s.energy()[0]
    s.step(int(round(T / h)))
    s.forward()
    return abs(s.energy()[0] - e0)


@pytest.mark.parametrize("env", ["HalfCheetah-v5", "Hopper-v5"])
def test_energy_conservation_order_planar(env):
    e = [_drift(env, 0, h) for h in (0.02, 0.01)]
    assert 1.8 < e[0] / e[1] < 2.2  # Euler: first order
    r = [_drift(env, 1, h) for h in (0.04, 0.02)]
    assert r[0] / r[1] > 12 and r[1] < 1e-4  # RK4: fourth order


@pytest.mark.parametrize("env", ["Ant-v5", "Humanoid-v4"])
def test_energy_conservation_free_joint(env):
    # quaternion stages are exponential-map updates from q0, so the scheme is second order on SO(3)
    e = [_drift(env, 1, h, T=0.2, vel=0.5) for h in (0.004, 0.002)]
    assert 3.5 < e[0] / e[1] < 4.5
    assert e[1] < 1e-5 * 500


@pytest.mark.parametrize("env", ["Ant-v5", "Humanoid-v4"])
def test_momentum_conserved_under_internal_actuation(env):
    s = Sim(env, flags=NO_CONTACT | NO_LIMIT | NO_GRAVITY, integrator=1, timestep=0.002)
    r = np.random.default_rng(3)
    s.qpos[:] = s.qpos + r.uniform(-0.2, 0.2, s.nq)
    s.qvel[:] = r.normal(size=s.nv) * 0.5
    s.ctrl[:] = r.uniform(-0.4, 0.4, s.nu)

    def momentum():
        s.forward()
        ci, cv, tot = s.cinert.reshape(-1, 10), s.cvel.reshape(-1, 6), np.zeros(6)
        for b in range(1, s.nbody):
            i, v = ci[b], cv[b]
            I = np.array([[i[0], i[3], i[4]], [i[3], i[1], i[5]], [i[4], i[5], i[2]]])
            tot[:3] += I @ v[:3] + np.cross(i[6:9], v[3:])
            tot[3:] += i[9] * v[3:] - np.cross(i[6:9], v[:3])
        return tot

    m0 = momentum()
    s.step(100)
    np.testing.assert_allclose(momentum(), m0, atol=2e-4 * max(1.0, np.abs(m0).max()))


def test_free_fall_closed_form():
    s = Sim("Ant-v5", flags=NO_CONTACT | NO_LIMIT)
    z0 = s.qpos[2]
    s.step(20)
    t = 20 * 0.01
    assert abs(s.qpos[2] - (z0 - 0.5 * 9.81 * t * t)) < 1e-12
    assert abs(s.qvel[2] + 9.81 * t) < 1e-12


@pytest.mark.parametrize("env,weight", [("HalfCheetah-v5", 14.0 * 9.81), ("Ant-v5", 0.9108800827 * 9.81)])
def test_resting_contact_force_equals_weight(env, weight):
    s = Sim(env)
    s.step(400)
    s.rne_post()
    fz = s.cfrc_ext.reshape(-1, 6)[:, 5].sum()
    assert abs(fz - weight) / weight < 2e-3
    assert np.abs(s.qvel).max() < 0.05


def test_newton_solution_satisfies_kkt():
    """At the solver's answer: M(qacc - qacc_smooth) = J^T f, f = -D*min(0, J qacc - aref)."""
    s = Sim("Ant-v5")
    s.step(150)
    s.forward()
    n = s.counts()["nefc"]
    assert n > 0
    J = s.efc_J.reshape(n, 24)[:, : s.nv]
    jar = J @ s.qacc - s.efc_aref
    f = np.where(jar < 0, -jar / s.efc_R, 0.0)
    np.testing.assert_allclose(s.efc_force, f, atol=1e-7 * max(1.0, np.abs(f).max()))
    lhs = s.mass_matrix() @ (s.qacc - s.qacc_smooth)
    np.testing.assert_allclose(lhs, J.T @ f, atol=1e-6 * max(1.0, np.abs(lhs).max()))


def test_pgs_forces_nonnegative_and_complementary():
    s = Sim("Humanoid-v4")
    s.step(300)
    s.forward()
    n = s.counts()["nefc"]
    assert n > 0 and np.all(s.efc_force >= 0)
    J = s.efc_J.reshape(n, 24)[:, : s.nv]
    lhs = s.mass_matrix() @ (s.qacc - s.qacc_smooth)
    np.testing.assert_allclose(lhs, J.T @ s.efc_force, atol=1e-8 * max(1.0, np.abs(lhs).max()))
