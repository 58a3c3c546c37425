"""CPU tests pinning the oracle's physics restatement with checks that do not need MuJoCo: algebraic identities of
the pipeline (M qacc = f, bias = dPE/dq at rest), contact force balance at rest, mj_stepSkip stage-reuse equivalence
(the property the device's no-skip schedule relies on), and Differentiator::DynamicsDerivatives against an
independent Python finite difference through Data.step.  MuJoCo itself is not installable here: parity with the real
library stays UNPINNED, these tests pin internal consistency."""
import numpy as np
import pytest

from helpers import make, oracle_model, oracle_task, perturbed_start, random_controls
from oracle import oracle
from trajoptkp_b200 import capi


def _snap_parts(m, s):
    q = s[1:1 + m.nq]
    v = s[1 + m.nq:1 + m.nq + m.nv]
    return q, v


@pytest.mark.parametrize("name", ["acrobot", "walker_run"])
def test_equation_of_motion_identity(built, name):
    tc = make(name)
    om = oracle_model(tc)
    d = oracle.Data(om)
    m = tc.model
    for seed in range(4):
        s = perturbed_start(tc, seed, dq=0.3, dv=1.0)
        s[m.S - m.nu:] = np.random.default_rng(seed).uniform(-1, 1, m.nu)
        d.load(s)
        d.forward()
        M = d.get("qM").reshape(m.nv, m.nv)
        assert np.allclose(M, M.T, atol=1e-13)
        assert np.linalg.eigvalsh(M).min() > 0
        rhs = d.get("qfrc_smooth") + d.get("qfrc_constraint")
        assert np.allclose(M @ d.get("qacc"), rhs, rtol=1e-9, atol=1e-9)
        assert np.allclose(M @ d.get("qacc_smooth"), d.get("qfrc_smooth"), rtol=1e-10, atol=1e-10)


@pytest.mark.parametrize("name", ["acrobot", "walker_run"])
def test_bias_force_is_potential_gradient_at_rest(built, name):
    """With qvel = 0, qfrc_bias = d/dq sum_b m_b g z_b (RNE against kinematics, independent paths)."""
    tc = make(name)
    om = oracle_model(tc)
    d = oracle.Data(om)
    m = tc.model
    mass = m.real("body_mass")

    def pe(q):
        s = tc.model.snapshot(qpos=q, qvel=np.zeros(m.nv))
        d.load(s)
        d.forward()
        return 9.81 * float(np.dot(mass, d.get("xipos").reshape(-1, 3)[:, 2]))

    rng = np.random.default_rng(3)
    q0 = m.real("qpos0") + rng.uniform(-0.4, 0.4, m.nq)
    if name == "walker_run":
        q0[0] += 1.0  # lift clear of the floor
    s = tc.model.snapshot(qpos=q0, qvel=np.zeros(m.nv))
    d.load(s)
    d.forward()
    bias = d.get("qfrc_bias")
    h = 1e-6
    grad = np.zeros(m.nv)
    for i in range(m.nv):  # hinge / slide only: nq == nv
        e = np.zeros(m.nq)
        e[i] = h
        grad[i] = (pe(q0 + e) - pe(q0 - e)) / (2 * h)
    assert np.allclose(bias, grad, rtol=1e-6, atol=1e-6)


def test_walker_contact_forces_balance_weight_at_rest(built):
    tc = make("walker_run")
    om = oracle_model(tc)
    d = oracle.Data(om)
    m = tc.model
    d.load(tc.start_snapshot())
    for _ in range(1500):
        d.step()
    v = d.get("qvel")
    assert np.abs(v).max() < 5e-2
    assert d.get("ncon")[0] >= 2
    total_weight = 9.81 * m.real("body_mass").sum()
    fz = d.get("qfrc_constraint")[0] if False else None  # rootz is dof 0 in walker_plane.xml
    fcon = d.get("qfrc_constraint")
    # vertical slide is the first dof (rootz): constraint force supports the weight (limits add nothing on rootz)
    assert abs(fcon[0] - total_weight) < 0.05 * total_weight, (fcon[0], total_weight)


@pytest.mark.parametrize("name", ["acrobot", "walker_run"])
def test_step_skip_equals_full_step(built, name):
    """mj_stepSkip(POS / VEL) after a nominal forward pass equals a full step from the perturbed snapshot when only
    velocities / controls changed - the fact that lets the device run every FD world as a full step
    (Differentiator.cpp:118,253,357 use the skip stages purely as an optimisation)."""
    tc = make(name)
    om = oracle_model(tc)
    m = tc.model
    a, b = oracle.Data(om), oracle.Data(om)
    for seed in range(3):
        s = perturbed_start(tc, 10 + seed, dq=0.05, dv=0.5)
        if name == "walker_run":
            for _ in range(0):
                pass
        s[m.S - m.nu:] = np.random.default_rng(seed).uniform(-0.5, 0.5, m.nu)
        # velocity perturbation, skip POS
        a.load(s)
        a.step(oracle.STAGE_NONE, 5, 0.0)          # nominal step leaves position-stage results of s in `a`
        sv = s.copy()
        sv[1 + m.nq + 1 % m.nv] += 1e-6
        a.load(sv)
        a.step(oracle.STAGE_POS, 5, 0.0)
        b.load(sv)
        b.step(oracle.STAGE_NONE, 5, 0.0)
        assert np.array_equal(a.save(), b.save())
        # control perturbation, skip VEL
        a.load(s)
        a.step(oracle.STAGE_NONE, 5, 0.0)
        su = s.copy()
        su[m.S - m.nu] += 1e-6
        a.load(su)
        a.step(oracle.STAGE_VEL, 5, 0.0)
        b.load(su)
        b.step(oracle.STAGE_NONE, 5, 0.0)
        assert np.array_equal(a.save(), b.save())


def test_integrate_differentiate_roundtrip(built):
    tc = make("walker_run")
    om = oracle_model(tc)
    m = tc.model
    rng = np.random.default_rng(0)
    q1 = m.real("qpos0") + rng.uniform(-0.3, 0.3, m.nq)
    v = rng.standard_normal(m.nv)
    q2 = q1.copy()
    oracle.lib().ok_integrate_pos(om.h, oracle._d(q2), oracle._d(v), 0.01)
    back = np.zeros(m.nv)
    oracle.lib().ok_differentiate_pos(om.h, oracle._d(back), 0.01, oracle._d(q1), oracle._d(q2))
    assert np.allclose(back, v, rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("name,tol", [("acrobot", 1e-9), ("walker_run", 1e-9)])
def test_dynamics_derivatives_match_python_fd(built, name, tol):
    """Differentiator::DynamicsDerivatives restated in Python on top of Data.step: same perturbations, same order
    (central, eps = 1e-6, FD solver options 5 / 0) must reproduce the oracle's A and B to rounding."""
    tc = make(name)
    om = oracle_model(tc)
    ot = oracle_task(tc)
    m = tc.model
    n, mm = tc.n, tc.m
    s = perturbed_start(tc, 42, dq=0.05, dv=0.3)
    s[m.S - m.nu:] = 0.3
    A, B = oracle.dynamics_derivs(om, ot, s, list(range(n)))
    d = oracle.Data(om)
    eps = 1e-6

    def next_state(snap):
        d.load(snap)
        d.step(oracle.STAGE_NONE, 5, 0.0)
        o = d.save()
        return o[1:1 + m.nq][tc.qpos_adr], o[1 + m.nq:1 + m.nq + m.nv][tc.dof_adr]

    A2 = np.zeros_like(A)
    B2 = np.zeros_like(B)
    for c in range(n):
        for kind in ("pos", "vel"):
            sp, sm = s.copy(), s.copy()
            idx = 1 + tc.qpos_adr[c] if kind == "pos" else 1 + m.nq + tc.dof_adr[c]
            sp[idx] += eps
            sm[idx] -= eps
            qp, vp = next_state(sp)
            qm, vm = next_state(sm)
            col = c if kind == "pos" else c + n
            A2[:n, col] = (qp - qm) / (2 * eps)
            A2[n:, col] = (vp - vm) / (2 * eps)
    for c in range(mm):
        sp, sm = s.copy(), s.copy()
        idx = m.S - m.nu + tc.ctrl_adr[c]
        sp[idx] += eps
        sm[idx] -= eps
        qp, vp = next_state(sp)
        qm, vm = next_state(sm)
        B2[:n, c] = (qp - qm) / (2 * eps)
        B2[n:, c] = (vp - vm) / (2 * eps)
    assert np.linalg.norm(A - A2) <= tol * max(1.0, np.linalg.norm(A2))
    assert np.linalg.norm(B - B2) <= tol * max(1.0, np.linalg.norm(B2))


def test_acrobot_ilqr_cost_decreases(built):
    """End-to-end oracle iLQR on the reference's CPU-runnable config (acrobot swing-up, set_interval key-points)."""
    tc = make("acrobot", kp_method=capi.KP_SET_INTERVAL, min_N=5)
    om = oracle_model(tc)
    ot = oracle_task(tc)
    T = 120
    il = oracle.ILQR(om, ot, T, threads=2)
    il.set_lambda(0.1)
    U0 = np.zeros((T, tc.m))
    il.optimise(tc.start_snapshot(), U0, 4, 2)
    sc = il.scalars()
    assert sc["new_cost"] <= sc["initial_cost"]
    assert sc["cost_reduction"] >= 0
    rows = il.keypoints()
    assert np.allclose(il.get("last_percentages"), 100.0 * sum(1 for r in rows if r) / T)
