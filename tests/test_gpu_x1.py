"""GPU: every BASELINE config on ITS OWN key-point method and at the bench's batch size:
  * box_sweep and push_lcl with iterative_error (key-point rows bit-exact vs the oracle, A / B <= 1e-4 rel. Frobenius),
    on trajectories where the gripper really pushes the object (the reference's Jacobian end-effector initial controls);
  * push_mcl: key-points, Jacobians and the optimised cost (it only had a step test);
  * one batch-size test per Franka config at bench.py's B with oracle spot checks and batch properties;
  * the quaternion branches of the FD code (mj_integratePos / mj_differentiatePos on a free joint's angular DoF).
Tolerances (north_star): key-point indices bit-exact; states 1e-9 (1e-8 on contact-rich scenes: converged Newton on both
sides, other summation order); A / B 1e-4 relative Frobenius; final cost within 1 %."""
import numpy as np
import pytest

from helpers import make, oracle_model, oracle_task, random_controls, rel_fro
from oracle import oracle
from trajoptkp_b200 import capi

pytestmark = pytest.mark.gpu

AB_RTOL = 1e-4


def _n2(tc, T, trials, setup):
    snaps, goals = zip(*[tc.init_snapshot(i) for i in trials])
    targets = [tc.load_trial(i)[1] for i in trials]
    tr = capi.Traj(tc, T, len(trials))
    for b in range(len(trials)):
        tr.set_targets(b, targets[b])
    snap0, U, st = tr.init_controls_ee(np.stack(snaps), np.stack(goals), pre_steps=1, setup_horizon=setup)
    assert not st.any(), st
    return tr, snap0, U, targets


def _compare_ab(tr, ref, rows, b, T):
    nx = tr.nx
    A_ref, B_ref = ref.get("A").reshape(T, nx, nx), ref.get("B").reshape(T, tr.m, nx)
    A_dev, B_dev = tr.get(b, "A").reshape(T, nx, nx), tr.get(b, "B").reshape(T, tr.m, nx)
    n = tr.n
    worst = 0.0
    for t in range(T):
        for j in set(rows[t]):   # only the columns this method differentiated at t
            cols = [j, j + n]
            worst = max(worst, rel_fro(A_dev[t][cols], A_ref[t][cols]))
            if j < tr.m:
                worst = max(worst, rel_fro(B_dev[t][j], B_ref[t][j]))
    return worst


@pytest.mark.parametrize("name,T,setup,min_N,thr", [("box_sweep", 60, 0, 1, 1e-5), ("push_lcl", 40, 120, 1, 1e-3)])
def test_iterative_error_on_the_configs_that_use_it(built, name, T, setup, min_N, thr):
    """BASELINE configs[2] / [3]: iterative_error with the thresholds of the task YAML / the published runs
    (box_sweep 1e-5, push_lcl 1e-3; KeyPointGenerator.cpp:449-640)."""
    tc = make(name, kp_method=capi.KP_ITERATIVE_ERROR, min_N=min_N, max_N=50)
    tc.iter_thr = thr
    tr, snap0, U, targets = _n2(tc, T, [0, 4], setup)
    tr.set_keypoint_method(capi.KP_ITERATIVE_ERROR, min_N, 50, iter_thr=thr)
    tr.rollout(snap0, U)
    rows, pct = tr.keypoints()   # iterative_error computes the Jacobian columns while it bisects
    om = oracle_model(tc)
    for b in range(2):
        ref = oracle.ILQR(om, oracle_task(tc, target=targets[b]), T, threads=1)
        ref.rollout(snap0[b], U[b])
        ref.generate_keypoints()
        assert rows[b] == ref.keypoints(), (name, b)
        total = sum(len(r) for r in rows[b])
        assert 3 * tc.n <= total < T * tc.n      # start, middle, end at least; not everything
        assert _compare_ab(tr, ref, rows[b], b, T) < AB_RTOL


def test_push_mcl_keypoints_jacobians_cost(built):
    """BASELINE configs[3], heavy clutter (nv = 63, 31 optimiser DoFs, set_interval as its YAML ships)."""
    tc = make("push_mcl", kp_method=capi.KP_SET_INTERVAL, min_N=4)
    T = 12
    tr, snap0, U, targets = _n2(tc, T, [1], 60)
    om = oracle_model(tc)
    ref = oracle.ILQR(om, oracle_task(tc, target=targets[0]), T, threads=8)
    c_ref = ref.rollout(snap0[0], U[0])
    ref.generate_keypoints()
    ref.dynamics_derivs()
    c_dev = tr.rollout(snap0, U)[0]
    assert abs(c_dev - c_ref) <= 1e-8 * max(1.0, abs(c_ref)), (c_dev, c_ref)
    X_ref, X_dev = ref.get("X_old"), tr.get(0, "X_old")
    assert np.abs(X_ref - X_dev).max() / max(1.0, np.abs(X_ref).max()) < 1e-8
    assert np.allclose(ref.get("resid"), tr.get(0, "resid"), rtol=0, atol=1e-8)
    rows, pct = tr.keypoints()
    assert rows[0] == ref.keypoints()
    tr.dynamics_derivs()
    assert _compare_ab(tr, ref, rows[0], 0, T) < AB_RTOL
    ref.set_lambda(0.1)
    ref.optimise(snap0[0], U[0], 2, 1)
    sc = ref.scalars()
    _, stats = tr.optimise(snap0, U, 2, 1, lam=0.1)
    assert abs(stats[0, 0] - sc["initial_cost"]) <= 1e-8 * abs(sc["initial_cost"])
    assert abs(stats[0, 1] - sc["new_cost"]) <= 0.01 * abs(sc["new_cost"]), (stats[0], sc)


@pytest.mark.parametrize("name,B,T,kp,setup", [("reaching", 512, 50, capi.KP_ADAPTIVE_JERK, 0),
                                               ("box_sweep", 296, 100, capi.KP_ITERATIVE_ERROR, 0),
                                               ("push_ncl", 256, 50, capi.KP_ADAPTIVE_JERK, 1000),
                                               ("push_lcl", 64, 100, capi.KP_ITERATIVE_ERROR, 1000),
                                               ("push_mcl", 16, 50, capi.KP_SET_INTERVAL, 1000)])
def test_bench_batch_sizes_properties_and_spot_checks(built, name, B, T, kp, setup):
    """The bench's own batch (bench.py DEFAULTS) per Franka config: one replan of every trajectory; batch slots are
    independent and deterministic (a duplicated problem gives identical bits), nothing overflows or warns, the cost
    never rises, and one slot is checked against the oracle end to end (initial cost 1e-8, final cost 1 %)."""
    tc = make(name, kp_method=kp)
    trials = [b % 100 for b in range(B)]
    trials[B - 1] = trials[0]
    if tc.kind == capi.TASK_REACHING:
        snaps = np.stack([tc.init_snapshot(i)[0] for i in trials])
        tr = capi.Traj(tc, T, B)
        targets = [tc.load_trial(i)[1] for i in trials]
        for b in range(B):
            tr.set_targets(b, targets[b])
        snap0 = snaps
        U, st = tr.init_controls_gravcomp(snaps)
        assert not st.any()
    else:
        tr, snap0, U, targets = _n2(tc, T, trials, setup)
    Uo, stats = tr.optimise(snap0, U, 1, 1, lam=0.1)
    assert np.isfinite(stats).all() and np.isfinite(Uo).all()
    assert not tr.status().any(), np.unique(tr.status())
    assert (stats[:, 1] <= stats[:, 0] * (1 + 1e-12)).all()
    assert np.array_equal(Uo[0], Uo[B - 1]) and np.array_equal(stats[0, :4], stats[B - 1, :4])
    assert np.array_equal(tr.get(0, "A"), tr.get(B - 1, "A"))
    # spot check against the oracle (slot 1; push_mcl's CPU replan takes a few seconds)
    b = 1
    ref = oracle.ILQR(oracle_model(tc), oracle_task(tc, target=targets[b]), T, threads=8)
    ref.set_lambda(0.1)
    ref.optimise(snap0[b], U[b], 1, 1)
    sc = ref.scalars()
    assert abs(stats[b, 0] - sc["initial_cost"]) <= 1e-8 * abs(sc["initial_cost"]), (stats[b], sc)
    assert abs(stats[b, 1] - sc["new_cost"]) <= 0.01 * abs(sc["new_cost"]), (stats[b], sc)


def test_angular_free_joint_dof_runs_the_quaternion_branches(built):
    """cuda_fd.cu: mj_integratePos on a free joint's angular DoF (quaternion exponential) and mj_differentiatePos
    (quaternion logarithm) - no shipped config activates an angular body DoF, so the task tables are edited here: the
    box's z position element is re-pointed at its rotation about z (dof address + 5).  Device vs oracle on every column."""
    tc = make("box_sweep", kp_method=capi.KP_SET_INTERVAL, min_N=3)
    m = tc.model
    jid = int(m.int("body_jntadr")[m.name2id(capi.OBJ_BODY, "goal")])
    da, qa = int(m.int("jnt_dofadr")[jid]), int(m.int("jnt_qposadr")[jid])
    tc.dof_adr = tc.dof_adr.copy()
    tc.qpos_adr = tc.qpos_adr.copy()
    tc.dof_adr[9] = da + 5          # rotation about the body z axis
    tc.qpos_adr[9] = qa + 6         # a quaternion component stands in for the "position" (axis-angle read-out is out of scope)
    T = 12
    snap, goal = tc.init_snapshot(3)
    f = m.split(snap)
    f["qpos"][qa + 3:qa + 7] = [np.cos(0.2), 0.0, 0.0, np.sin(0.2)]   # a yawed box, so that q is not the identity
    f["qvel"][da + 5] = 0.3
    snap0 = m.snapshot(qpos=f["qpos"], qvel=f["qvel"])
    U = random_controls(tc, T, 21, scale=0.1)
    tr = capi.Traj(tc, T, 1)
    tr.rollout(snap0, U)
    rows, _ = tr.keypoints()
    tr.dynamics_derivs()
    ref = oracle.ILQR(oracle_model(tc), oracle_task(tc), T, threads=4)
    ref.rollout(snap0, U)
    ref.generate_keypoints()
    ref.dynamics_derivs()
    assert rows[0] == ref.keypoints()
    nx = tr.nx
    A_ref, A_dev = ref.get("A").reshape(T, nx, nx), tr.get(0, "A").reshape(T, nx, nx)
    kp = [t for t in range(T) if rows[0][t]]
    assert max(rel_fro(A_dev[t], A_ref[t]) for t in kp) < AB_RTOL
    # the angular column and row are not trivially zero: d(yaw')/d(yaw) ~ 1, d(yaw')/d(yaw rate) ~ dt (floor friction
    # acts on the spinning box, so not exactly)
    for t in kp:
        assert abs(A_dev[t][9][9] - 1.0) < 0.05, A_dev[t][9][9]              # column-major block: [col][row]
        assert abs(A_dev[t][19][9] - m.real("opt_real")[0]) < 0.5 * m.real("opt_real")[0]
