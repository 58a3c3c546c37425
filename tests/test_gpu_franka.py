"""GPU parity on the Panda reaching config (BASELINE configs[1]; reaching_scene.xml: nq 16 / nv 15 / 13 bodies, 11 mesh
collision geoms, finger pads in permanent mesh-mesh contact): device path through the C ABI against the CPU oracle on
the same seeded inputs.  Tolerances as in test_gpu_parity.py (north_star): per-step states 1e-9 relative, A/B 1e-4
relative Frobenius (contacts), key-point rows bit-exact, final cost within 1 %."""
import numpy as np
import pytest

from helpers import make, oracle_model, oracle_task, perturbed_start, random_controls, rel_fro
from oracle import oracle
from trajoptkp_b200 import capi

pytestmark = pytest.mark.gpu

STATE_RTOL = 1e-9
AB_RTOL = 1e-4


def test_reaching_step_worlds_match_oracle(built):
    tc = make("reaching")
    m = tc.model
    om = oracle_model(tc)
    rng = np.random.default_rng(5)
    count, nsteps = 16, 10
    snaps = np.stack([perturbed_start(tc, 300 + i, dq=0.3, dv=0.5) for i in range(count)])
    lim = m.real("actuator_ctrlrange").reshape(-1, 2)[:, 1]
    snaps[:, m.S - m.nu:] = rng.uniform(-0.5, 0.5, (count, m.nu)) * lim
    out = np.zeros_like(snaps)
    st = np.zeros(count, dtype=np.int32)
    capi.check(capi.lib().tokp_step_worlds(m.h, capi._d(snaps), capi._d(out), count, nsteps, -1, 0.0, capi._i(st)))
    assert not st.any(), st
    d = oracle.Data(om)
    contacts = 0
    for i in range(count):
        d.load(snaps[i])
        for _ in range(nsteps):
            d.step()
        assert d.get("warning")[0] == 0
        contacts += int(d.get("ncon")[0])
        ref = d.save()
        err = np.abs(ref - out[i]).max() / max(1.0, np.abs(ref).max())
        assert err < STATE_RTOL, (i, err)
    assert contacts > 0  # the mesh-mesh collider is exercised (finger pads)


def test_reaching_rollout_keypoints_jacobians(built):
    tc = make("reaching", kp_method=capi.KP_ADAPTIVE_JERK, min_N=1, max_N=10)
    T = 30
    om, ot = oracle_model(tc), oracle_task(tc)
    snap0 = perturbed_start(tc, 9)
    U = random_controls(tc, T, 4, scale=0.2)
    ref = oracle.ILQR(om, ot, T, threads=4)
    c_ref = ref.rollout(snap0, U)
    ref.generate_keypoints()
    ref.dynamics_derivs()
    tr = capi.Traj(tc, T, 1)
    c_dev = tr.rollout(snap0, U)[0]
    assert abs(c_dev - c_ref) <= 1e-9 * max(1.0, abs(c_ref))
    X_ref, X_dev = ref.get("X_old"), tr.get(0, "X_old")
    assert np.abs(X_ref - X_dev).max() / max(1.0, np.abs(X_ref).max()) < STATE_RTOL
    rows, pct = tr.keypoints()
    assert rows[0] == ref.keypoints()  # adaptive_jerk rows bit-exact
    tr.dynamics_derivs()
    nx = tr.nx
    A_ref, B_ref = ref.get("A").reshape(T, nx, nx), ref.get("B").reshape(T, tr.m, nx)
    A_dev, B_dev = tr.get(0, "A").reshape(T, nx, nx), tr.get(0, "B").reshape(T, tr.m, nx)
    kp_rows = [t for t in range(T) if rows[0][t]]
    worst = max(max(rel_fro(A_dev[t], A_ref[t]), rel_fro(B_dev[t], B_ref[t])) for t in kp_rows)
    assert worst < AB_RTOL, worst


def test_reaching_optimise_cost(built):
    tc = make("reaching", kp_method=capi.KP_SET_INTERVAL, min_N=5)
    T = 40
    om = oracle_model(tc)
    start, target = tc.load_trial(3)
    ot = oracle_task(tc, target=target)
    snap0 = tc.start_snapshot(start_pos=start)
    U = np.zeros((T, tc.m))
    ref = oracle.ILQR(om, ot, T, threads=4)
    ref.set_lambda(0.1)
    ref.optimise(snap0, U, 3, 1)
    sc = ref.scalars()
    tr = capi.Traj(tc, T, 2)
    for b in range(2):
        tr.set_targets(b, target)
    U_dev, stats = tr.optimise(snap0, U, 3, 1, lam=0.1)
    for b in range(2):
        assert abs(stats[b, 0] - sc["initial_cost"]) <= 1e-9 * abs(sc["initial_cost"])
        assert abs(stats[b, 1] - sc["new_cost"]) <= 0.01 * abs(sc["new_cost"]), (stats[b], sc)  # final cost within 1 %
        assert int(stats[b, 2]) == int(sc["num_iterations"])
    assert stats[0, 1] < stats[0, 0]


def test_reaching_gravcomp_init_controls(built):
    """tokp_init_controls_gravcomp == pandaReaching::CreateInitOptimisationControls restated with the oracle:
    mj_forward, ctrl := qfrc_bias of the arm joints, mj_step (Reaching.cpp:269-296)."""
    tc = make("reaching")
    m = tc.model
    T, B = 25, 3
    om = oracle_model(tc)
    snaps = np.stack([perturbed_start(tc, 40 + b, dq=0.1, dv=0.05) for b in range(B)])
    tr = capi.Traj(tc, T, B)
    U_dev, st = tr.init_controls_gravcomp(snaps)
    assert not st.any()
    d = oracle.Data(om)
    for b in range(B):
        d.load(snaps[b])
        for t in range(T):
            d.forward()
            bias = d.get("qfrc_bias")
            s = d.save()
            s[m.S - m.nu + tc.ctrl_adr] = bias[tc.dof_adr]
            d.load(s)
            d.step()
            ref = bias[tc.dof_adr]
            assert np.abs(U_dev[b, t] - ref).max() <= 1e-9 * max(1.0, np.abs(ref).max()), (b, t)
    # holding the arm: joint velocities stay small under gravity compensation
    assert np.abs(U_dev).max() > 1.0


# ---- box_sweep (BASELINE configs[2]): boxSweep.xml nq 23 / nv 21, state = 7 arm joints + goal box x, y, z; contacts:
# box-plane (condim 1), mesh-box through the convex collider, finger pads, the display_goal sliver on the floor

def test_box_sweep_step_worlds_match_oracle(built):
    tc = make("box_sweep")
    m = tc.model
    om = oracle_model(tc)
    rng = np.random.default_rng(8)
    count, nsteps = 12, 12
    snaps = []
    for i in range(count):
        start, _ = tc.load_trial(i)
        r = np.random.default_rng(500 + i)
        start[:7] += r.uniform(-0.05, 0.05, 7)
        start[9] = 0.081 + 0.05 * (i % 3)   # box resting on / dropping onto the floor
        snaps.append(tc.start_snapshot(start_pos=start, qvel=0.1 * r.standard_normal(tc.n)))
    snaps = np.stack(snaps)
    lim = m.real("actuator_ctrlrange").reshape(-1, 2)[:, 1]
    snaps[:, m.S - m.nu:] = rng.uniform(-0.3, 0.3, (count, m.nu)) * lim
    out = np.zeros_like(snaps)
    st = np.zeros(count, dtype=np.int32)
    capi.check(capi.lib().tokp_step_worlds(m.h, capi._d(snaps), capi._d(out), count, nsteps, -1, 0.0, capi._i(st)))
    assert not st.any(), st
    d = oracle.Data(om)
    worst, ncon = 0.0, 0
    for i in range(count):
        d.load(snaps[i])
        for _ in range(nsteps):
            d.step()
        assert d.get("warning")[0] == 0
        ncon = max(ncon, int(d.get("ncon")[0]))
        ref = d.save()
        worst = max(worst, np.abs(ref - out[i]).max() / max(1.0, np.abs(ref).max()))
    assert ncon >= 9   # box on the floor (4) + display_goal (4) + finger pads
    # contact-rich: converged Newton on both sides; summation order differs
    assert worst < 1e-8, worst


def test_box_sweep_rollout_keypoints_jacobians(built):
    tc = make("box_sweep", kp_method=capi.KP_SET_INTERVAL, min_N=4)
    T = 24
    om = oracle_model(tc)
    start, target = tc.load_trial(1)
    start[9] = 0.0805
    ot = oracle_task(tc, target=target)
    snap0 = tc.start_snapshot(start_pos=start)
    U = random_controls(tc, T, 6, scale=0.1)
    ref = oracle.ILQR(om, ot, T, threads=4)
    c_ref = ref.rollout(snap0, U)
    ref.generate_keypoints()
    ref.dynamics_derivs()
    tr = capi.Traj(tc, T, 1)
    tr.set_targets(0, target)
    c_dev = tr.rollout(snap0, U)[0]
    assert abs(c_dev - c_ref) <= 1e-8 * max(1.0, abs(c_ref)), (c_dev, c_ref)
    X_ref, X_dev = ref.get("X_old"), tr.get(0, "X_old")
    assert np.abs(X_ref - X_dev).max() / max(1.0, np.abs(X_ref).max()) < 1e-8
    assert np.allclose(ref.get("resid"), tr.get(0, "resid"), rtol=0, atol=1e-8)   # incl. the gripper-to-box distance
    rows, pct = tr.keypoints()
    assert rows[0] == ref.keypoints()
    tr.dynamics_derivs()
    nx = tr.nx
    A_ref, B_ref = ref.get("A").reshape(T, nx, nx), ref.get("B").reshape(T, tr.m, nx)
    A_dev, B_dev = tr.get(0, "A").reshape(T, nx, nx), tr.get(0, "B").reshape(T, tr.m, nx)
    kp_rows = [t for t in range(T) if rows[0][t]]
    errs = [max(rel_fro(A_dev[t], A_ref[t]), rel_fro(B_dev[t], B_ref[t])) for t in kp_rows]
    assert max(errs) < AB_RTOL, errs


def test_box_sweep_optimise_cost(built):
    tc = make("box_sweep", kp_method=capi.KP_SET_INTERVAL, min_N=5)
    T = 40
    om = oracle_model(tc)
    start, target = tc.load_trial(2)
    ot = oracle_task(tc, target=target)
    snap0 = tc.start_snapshot(start_pos=start)
    U = np.zeros((T, tc.m))
    ref = oracle.ILQR(om, ot, T, threads=4)
    ref.set_lambda(0.1)
    ref.optimise(snap0, U, 2, 1)
    sc = ref.scalars()
    tr = capi.Traj(tc, T, 2)
    for b in range(2):
        tr.set_targets(b, target)
    U_dev, stats = tr.optimise(snap0, U, 2, 1, lam=0.1)
    for b in range(2):
        assert abs(stats[b, 0] - sc["initial_cost"]) <= 1e-8 * abs(sc["initial_cost"])
        assert abs(stats[b, 1] - sc["new_cost"]) <= 0.01 * abs(sc["new_cost"]), (stats[b], sc)  # final cost within 1 %


# ---- push_ncl / push_lcl / push_mcl (BASELINE configs[3]): cylinder_pushing*.xml, nv 21 / 39 / 63; plane-cylinder,
# cylinder-cylinder / box-cylinder / mesh-cylinder through the convex collider, condim 1 and 3

def _push_snaps(tc, count, seed):
    snaps = []
    for i in range(count):
        start, _ = tc.load_trial(i)
        r = np.random.default_rng(seed + i)
        start[:7] += r.uniform(-0.05, 0.05, 7)
        snaps.append(tc.start_snapshot(start_pos=start, qvel=0.05 * r.standard_normal(tc.n)))
    return np.stack(snaps)


@pytest.mark.parametrize("name,count,nsteps", [("push_ncl", 12, 12), ("push_lcl", 8, 10), ("push_mcl", 4, 8)])
def test_push_step_worlds_match_oracle(built, name, count, nsteps):
    tc = make(name)
    m = tc.model
    om = oracle_model(tc)
    rng = np.random.default_rng(3)
    snaps = _push_snaps(tc, count, 700)
    lim = m.real("actuator_ctrlrange").reshape(-1, 2)[:, 1]
    snaps[:, m.S - m.nu:] = rng.uniform(-0.3, 0.3, (count, m.nu)) * lim
    out = np.zeros_like(snaps)
    st = np.zeros(count, dtype=np.int32)
    capi.check(capi.lib().tokp_step_worlds(m.h, capi._d(snaps), capi._d(out), count, nsteps, -1, 0.0, capi._i(st)))
    assert not st.any(), st
    d = oracle.Data(om)
    worst = 0.0
    for i in range(count):
        d.load(snaps[i])
        for _ in range(nsteps):
            d.step()
        assert d.get("warning")[0] == 0
        ref = d.save()
        worst = max(worst, np.abs(ref - out[i]).max() / max(1.0, np.abs(ref).max()))
    assert worst < 1e-8, worst


@pytest.mark.parametrize("name", ["push_ncl", "push_lcl"])
def test_push_rollout_keypoints_jacobians_cost(built, name):
    tc = make(name, kp_method=capi.KP_SET_INTERVAL, min_N=5)
    T = 20
    om = oracle_model(tc)
    start, target = tc.load_trial(1)
    ot = oracle_task(tc, target=target)
    snap0 = tc.start_snapshot(start_pos=start)
    U = random_controls(tc, T, 8, scale=0.1)
    ref = oracle.ILQR(om, ot, T, threads=4)
    c_ref = ref.rollout(snap0, U)
    ref.generate_keypoints()
    ref.dynamics_derivs()
    tr = capi.Traj(tc, T, 1)
    tr.set_targets(0, target)
    c_dev = tr.rollout(snap0, U)[0]
    assert abs(c_dev - c_ref) <= 1e-8 * max(1.0, abs(c_ref)), (c_dev, c_ref)
    assert np.allclose(ref.get("resid"), tr.get(0, "resid"), rtol=0, atol=1e-8)
    rows, pct = tr.keypoints()
    assert rows[0] == ref.keypoints()
    tr.dynamics_derivs()
    nx = tr.nx
    A_ref, B_ref = ref.get("A").reshape(T, nx, nx), ref.get("B").reshape(T, tr.m, nx)
    A_dev, B_dev = tr.get(0, "A").reshape(T, nx, nx), tr.get(0, "B").reshape(T, tr.m, nx)
    kp_rows = [t for t in range(T) if rows[0][t]]
    errs = [max(rel_fro(A_dev[t], A_ref[t]), rel_fro(B_dev[t], B_ref[t])) for t in kp_rows]
    assert max(errs) < AB_RTOL, errs
    # one Optimise call: final cost within 1 %
    ref.set_lambda(0.1)
    ref.optimise(snap0, np.zeros((T, tc.m)), 2, 1)
    sc = ref.scalars()
    U_dev, stats = tr.optimise(snap0, np.zeros((T, tc.m)), 2, 1, lam=0.1)
    assert abs(stats[0, 1] - sc["new_cost"]) <= 0.01 * abs(sc["new_cost"]), (stats[0], sc)
