"""Closed-form pins of the oracle's physics restatement (SURVEY.md 8c plan item 2; VERDICT r1 "Next round" 2b, 2d).

MuJoCo itself is not available here (parity to the library stays unpinned); these tests tie the oracle to
the library's DOCUMENTED formulas through results that can be written down independently of the code under test:

1. static penetration depth of a resting body from solref / solimp / invweight0 (one scalar equation, solved by bisection here);
2. semi-implicit Euler with implicit joint damping ("eulerdamp") against the exact discrete map of the linearised system and
   against the analytic damped-oscillator solution (first order in h);
3. dof_invweight0 / body_invweight0 / meaninertia of the generated tables against values measured from the ORACLE'S OWN
   kinematics and CRB mass matrix by finite differences (a code path that shares nothing with mjcf_compile._set_const).
"""
import numpy as np
import pytest

from oracle.pbo import NO_CONTACT, NO_GRAVITY, NO_LIMIT, Sim


def _impedance(solimp, r):
    d0, dw, width, mid, power = solimp
    x = abs(r) / width
    if x >= 1:
        return dw
    y = x ** power / mid ** (power - 1) if x <= mid else 1 - (1 - x) ** power / (1 - mid) ** (power - 1)
    return d0 + y * (dw - d0)


def _pyramid_normal_force(dist, cp, tran):
    """Normal force of ONE resting condim-3 pyramidal contact (velocity 0, acceleration 0) from the documented formulas:
    aref = -k imp (dist - margin); every one of the 4 edge rows carries f = -(-aref) / R_py ... = k imp |r| / R_py with
    R_first = (1 - imp) / imp * tran (1 + mu^2), R_py = 2 mu^2 R_first; the edges' tangential parts cancel, normal parts add."""
    r = dist - cp["includemargin"]
    if r >= 0:
        return 0.0
    timeconst, dampratio = cp["solref"]
    dmax, mu = cp["solimp"][1], cp["friction"][0]
    k = 1.0 / (dmax * dmax * timeconst * timeconst * dampratio * dampratio)
    imp = _impedance(cp["solimp"], r)
    r_first = (1 - imp) / imp * tran * (1 + mu * mu)
    return 4.0 * k * imp * (-r) / (2 * mu * mu * r_first)


def test_static_penetration_depth_matches_solref_solimp_closed_form():
    s = Sim("Ant-v5")
    s.step(1500)
    s.forward()
    assert np.abs(s.qvel).max() < 1e-3 and np.abs(s.qacc).max() < 1e-3     # at rest
    n = s.counts()["ncon"]
    assert n == 4
    weight = s.model_field("body_mass").sum() * 9.81
    binv = s.model_field("body_invweight0").reshape(-1, 2)
    gb = s.model_field("geom_body")
    total, dists = 0.0, []
    for i in range(n):
        c, cp = s.contact(i), s.contact_params(i)
        tran = binv[gb[c["g1"]], 0] + binv[gb[c["g2"]], 0]
        f_pred = _pyramid_normal_force(c["dist"], cp, tran)
        f_orc = s.efc_force[c["efc"]:c["efc"] + 4].sum()
        assert abs(f_pred - f_orc) < 2e-5 * weight, (i, f_pred, f_orc)        # per contact: solver's force = closed form at its depth
        total += f_pred
        dists.append(c["dist"])
    assert abs(total - weight) < 1e-4 * weight
    # and the depth itself: the four legs are symmetric, so 4 F(d) = m g is one scalar equation for the resting distance d
    cp, c0 = s.contact_params(0), s.contact(0)
    tran = binv[gb[c0["g1"]], 0] + binv[gb[c0["g2"]], 0]
    lo, hi = cp["includemargin"] - 0.005, cp["includemargin"]
    for _ in range(80):
        mid = 0.5 * (lo + hi)
        if 4 * _pyramid_normal_force(mid, cp, tran) > weight:
            lo = mid
        else:
            hi = mid
    assert abs(0.5 * (lo + hi) - np.mean(dists)) < 2e-7, (0.5 * (lo + hi), np.mean(dists))


def test_eulerdamp_matches_exact_discrete_map_and_analytic_oscillator():
    """HalfCheetah-v5: Euler integrator, joint stiffness + damping + armature (half_cheetah.xml:38,63-81).  For small motion
    about qpos0 without gravity / contacts / limits the system is M x'' + B x' + K x = 0 and one mj_step is
        v+ = v + h (M + h B)^-1 (-K x - B v),   x+ = x + h v+
    (implicit in the damping, explicit in the stiffness).  The oracle must follow that map to rounding, and the map must
    approach the analytic solution exp(A t) of the continuous system at first order in h."""
    from scipy.linalg import expm
    amp, T = 1e-6, 0.2
    errs = []
    for h in (0.01, 0.005):
        s = Sim("HalfCheetah-v5", flags=NO_GRAVITY | NO_CONTACT | NO_LIMIT, timestep=h)
        nv = s.nv
        q0 = s.model_field("qpos0")
        s.forward()
        M = s.mass_matrix()
        B = np.diag(s.model_field("dof_damping"))
        K = np.zeros((nv, nv))
        stiff, dofadr, ref = s.model_field("jnt_stiffness"), s.model_field("jnt_dofadr"), s.model_field("jnt_springref")
        for j, d in enumerate(dofadr):
            K[d, d] = stiff[j]
        assert np.allclose(ref, 0) and (np.diag(B) > 0).sum() >= 6 and (stiff > 0).sum() >= 6
        rng = np.random.default_rng(4)
        x = np.zeros(nv)
        x[3:] = rng.normal(size=nv - 3) * amp          # joint angles only: the root dofs carry no spring
        v = rng.normal(size=nv) * amp * 5
        s.qpos[:] = q0 + x
        s.qvel[:] = v
        n = int(round(T / h))
        A = np.linalg.inv(M + h * B)
        xm, vm = x.copy(), v.copy()
        for _ in range(n):
            s.step()
            vm = vm + h * (A @ (-K @ xm - B @ vm))
            xm = xm + h * vm
        xo, vo = s.qpos - q0, s.qvel.copy()
        # the map itself: M(q) varies at O(amp), so agreement is to amp^2-level, far below any explicit-damping variant (O(h))
        assert np.abs(xo - xm).max() < 1e-4 * amp and np.abs(vo - vm).max() < 1e-4 * amp * 5, (np.abs(xo - xm).max(), np.abs(vo - vm).max())
        Ac = np.block([[np.zeros((nv, nv)), np.eye(nv)], [-np.linalg.solve(M, K), -np.linalg.solve(M, B)]])
        za = expm(Ac * T) @ np.concatenate([x, v])
        errs.append(np.abs(np.concatenate([xo, vo]) - za).max())
        # an explicit-damping Euler step would be v+ = v + h M^-1 (-K x - B v): make sure the test can tell the two apart
        xe, ve = x.copy(), v.copy()
        Mi = np.linalg.inv(M)
        for _ in range(n):
            ve = ve + h * (Mi @ (-K @ xe - B @ ve))
            xe = xe + h * ve
        assert np.abs(xe - xm).max() > 1e-2 * amp
    assert 1.6 < errs[0] / errs[1] < 2.4, errs      # first-order convergence to the analytic damped oscillator


def _vee(S):
    return np.array([S[2, 1] - S[1, 2], S[0, 2] - S[2, 0], S[1, 0] - S[0, 1]]) * 0.5


@pytest.mark.parametrize("env", ["HalfCheetah-v5", "Hopper-v5", "Ant-v5", "Humanoid-v4"])
def test_invweight0_and_meaninertia_from_the_oracles_own_kinematics(env):
    """mj_setConst's quantities re-measured without mjcf_compile: M from the oracle's CRB at qpos0, body Jacobians from
    central differences of the oracle's kinematics (xipos, xmat) along every dof - free-joint rotations perturbed in the
    body frame, as mj_integratePos does - then invweight = mean diagonal of J M^-1 J^T."""
    s = Sim(env, flags=NO_CONTACT | NO_LIMIT)
    nv, nb = s.nv, s.nbody
    q0 = s.model_field("qpos0")
    jt, qadr, dadr = s.model_field("jnt_type"), s.model_field("jnt_qposadr"), s.model_field("jnt_dofadr")

    def kin(q):
        s.qpos[:] = q
        s.qvel[:] = 0
        s.forward()
        return s.xipos.reshape(-1, 3).copy(), s.xmat.reshape(-1, 3, 3).copy()

    x0, R0 = kin(q0)
    M = s.mass_matrix()
    Minv = np.linalg.inv(M)
    assert abs(np.trace(M) / nv - s.model_field("meaninertia")[0]) < 1e-9 * np.trace(M)
    eps = 1e-6
    J = np.zeros((nb, 6, nv))

    def perturbed(dof_q, sign):
        q = q0.copy()
        dof_q(q, sign * eps)
        return kin(q)

    for j in range(len(jt)):
        if jt[j] == 0:  # free: 3 world-frame translations, 3 body-frame rotations
            movers = []
            for k in range(3):
                movers.append((dadr[j] + k, lambda q, e, k=k, a=qadr[j]: q.__setitem__(a + k, q[a + k] + e)))
            for k in range(3):
                def rot(q, e, k=k, a=qadr[j]):
                    w, x, y, z = q[a + 3:a + 7]
                    d = np.zeros(4); d[0] = np.cos(e / 2); d[1 + k] = np.sin(e / 2)
                    q[a + 3:a + 7] = [w * d[0] - x * d[1] - y * d[2] - z * d[3], w * d[1] + x * d[0] + y * d[3] - z * d[2],
                                      w * d[2] - x * d[3] + y * d[0] + z * d[1], w * d[3] + x * d[2] - y * d[1] + z * d[0]]
                movers.append((dadr[j] + 3 + k, rot))
        else:
            movers = [(dadr[j], lambda q, e, a=qadr[j]: q.__setitem__(a, q[a] + e))]
        for dof, mv in movers:
            xp, Rp = perturbed(mv, +1)
            xm, Rm = perturbed(mv, -1)
            for b in range(1, nb):
                J[b, :3, dof] = (xp[b] - xm[b]) / (2 * eps)
                J[b, 3:, dof] = _vee((Rp[b] - Rm[b]) @ R0[b].T) / (2 * eps)
    dof_inv = np.diag(Minv).copy()
    for j in range(len(jt)):
        if jt[j] == 0:
            d = dadr[j]
            dof_inv[d:d + 3] = dof_inv[d:d + 3].mean()
            dof_inv[d + 3:d + 6] = dof_inv[d + 3:d + 6].mean()
    np.testing.assert_allclose(s.model_field("dof_invweight0"), dof_inv, rtol=1e-8)
    tab = s.model_field("body_invweight0").reshape(-1, 2)
    for b in range(1, nb):
        A = J[b] @ Minv @ J[b].T
        np.testing.assert_allclose(tab[b], [np.trace(A[:3, :3]) / 3, np.trace(A[3:, 3:]) / 3], rtol=2e-6, atol=1e-9)
