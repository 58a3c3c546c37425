"""GPU: the reference's call sequence driven through the C++ shim classes (Differentiator / KeypointGenerator /
iLQR mirrors in trajoptkp_b200/host/) gives the same numbers as the ctypes path and matches the oracle."""
import json
import math
import os
import subprocess

import numpy as np
import pytest

from helpers import make, oracle_model, oracle_task
from oracle import oracle
from trajoptkp_b200 import capi

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "trajoptkp_b200", "_build", "shim_test")
XML = os.path.join(ROOT, "assets", "models", "acrobot.xml")


def test_shim_call_sequence_matches_ctypes_and_oracle(built):
    p = subprocess.run([EXE, XML], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    out = json.loads(p.stdout)
    T = 50
    tc = make("acrobot", kp_method=capi.KP_SET_INTERVAL, min_N=5, max_N=50)
    snap0 = tc.start_snapshot(start_pos=np.array([0.3, -0.2]))
    U = np.array([[0.2 * math.sin(0.3 * t)] for t in range(T)])
    tr = capi.Traj(tc, T, 1)
    cost = tr.rollout(snap0, U)[0]
    assert out["rollout_cost"] == cost
    rows, pct = tr.keypoints()
    assert [len(r) for r in rows[0]] == [int(c) for c in out["keypoint_counts"]]
    assert np.array_equal(pct[0], out["percentages"])
    assert out["cols_vs_batched_maxabs"] == 0.0
    assert out["interpolation_identity_exact"] == 1
    # Differentiator::ResidualDerivatives through the shim == ctypes path at the same saved state
    tr.cost_derivs()
    rx5 = tr.get(0, "r_x").reshape(T + 1, 4, tc.nr)[5]          # [column][residual]
    ru5 = tr.get(0, "r_u").reshape(T + 1, 1, tc.nr)[5]
    assert np.array_equal(np.array(out["r_x5"]).reshape(tc.nr, 4), rx5.T)
    assert np.array_equal(np.array(out["r_u5"]), ru5[0])
    tr.generate_derivatives()
    A5 = tr.get(0, "A").reshape(T, 4, 4)[5]
    assert np.array_equal(A5.ravel(), np.array(out["A5"]))
    # Optimiser::FilterDynamicsMatrices through the shim == numpy restatement of the reference filters
    import filter_numpy as fnp
    assert out["filter_position_rows_untouched"] == 1
    assert np.array_equal(np.array(out["A_lowpass_21"]), fnp.low_pass(np.array(out["A_raw_21"])))
    assert np.array_equal(np.array(out["A_fir_32"]), fnp.fir(np.array(out["A_raw_32"])))
    assert np.array_equal(np.array(out["A_raw_21"]), tr.get(0, "A").reshape(T, 4, 4)[:, 1, 2])
    # oracle: Differentiator::DynamicsDerivatives at the same saved state
    om, ot = oracle_model(tc), oracle_task(tc)
    A_ref, _ = oracle.dynamics_derivs(om, ot, tr.get_snapshot(0, 5), [0, 1])
    assert np.linalg.norm(A5.T - A_ref) <= 1e-7 * np.linalg.norm(A_ref)
    Uo, stats = capi.Traj(tc, T, 1).optimise(snap0, U, 3, 1, lam=0.1)
    assert np.array_equal(Uo[0, :, 0], out["U_opt"])
    assert stats[0, 0] == out["initial_cost"] and stats[0, 1] == out["final_cost"]
    il = oracle.ILQR(om, ot, T, threads=2)
    il.set_lambda(0.1)
    il.optimise(snap0, U, 3, 1)
    assert abs(il.scalars()["new_cost"] - out["final_cost"]) <= 0.01 * abs(out["final_cost"])


def test_shim_drives_box_sweep_through_the_reference_classes(built, tmp_path):
    """the C++ boundary on a TARGET config.  shim_test --boxsweep builds a ModelTranslator with the
    seven arm joints + the goal body's x, y, z (rigid-body state elements), vector targets and aux bodies, and runs
    RolloutTrajectory -> SetCurrentKeypointMethod(iterative_error) -> GenerateKeyPoints -> GenerateDerivatives ->
    Differentiator::DynamicsDerivatives -> iLQR::Optimise on boxSweep.xml; every number must equal the ctypes path's."""
    T = 40
    tc = make("box_sweep", kp_method=capi.KP_ITERATIVE_ERROR, min_N=1, max_N=50)
    tc.iter_thr = 1e-5
    snap, goal = tc.init_snapshot(2)
    target = tc.load_trial(2)[1]
    tr = capi.Traj(tc, T, 1)
    tr.set_targets(0, target)
    snap0, U, st = tr.init_controls_ee(snap[None], goal[None], pre_steps=1, setup_horizon=0)
    assert not st.any()
    inp = tmp_path / "boxsweep_input.txt"
    with open(inp, "w") as f:
        f.write("%d\n" % T)
        f.write(" ".join("%.17g" % v for v in snap0[0]) + "\n")
        f.write("%.17g %.17g\n" % (goal[0], goal[1]))
        for t in range(T):
            f.write(" ".join("%.17g" % v for v in U[0, t]) + "\n")
    xml = os.path.join(ROOT, "assets", "models", "franka", "boxSweep.xml")
    p = subprocess.run([EXE, xml, "--boxsweep", str(inp)], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    out = json.loads(p.stdout)
    assert (out["dof"], out["num_ctrl"], out["num_resid"]) == (tc.n, tc.m, tc.nr) == (10, 7, 3)
    assert np.allclose(target, [goal[0], goal[1], 0.0, 0.0, 0.05])       # what the driver hard-codes from the YAML
    # ModelTranslator::ReturnStateVector with body elements: robot joints, then goal x, y, z; then the velocities
    f = tc.model.split(snap0[0])
    assert np.array_equal(out["start_state"], np.concatenate([f["qpos"][tc.qpos_adr], f["qvel"][tc.dof_adr]]))
    tr.set_keypoint_method(capi.KP_ITERATIVE_ERROR, 1, 50, jerk_thr=tc.jerk_thr, vel_thr=tc.vel_thr, iter_thr=1e-5)
    cost = tr.rollout(snap0, U)[0]
    assert out["rollout_cost"] == cost
    # host BoxSweep::Residuals at the start state == the oracle's residuals there
    om, ot = oracle_model(tc), oracle_task(tc, target=target)
    r_ref = np.zeros(tc.nr)
    import ctypes as C
    d = ot.desc()
    assert oracle.lib().ok_residuals(om.h, C.byref(d), oracle._d(np.ascontiguousarray(snap0[0])), oracle._d(r_ref)) == 0
    assert np.allclose(out["start_residuals"], r_ref, rtol=0, atol=1e-12)
    rows, pct = tr.keypoints()
    assert [len(r) for r in rows[0]] == [int(c) for c in out["keypoint_counts"]]
    assert np.array_equal(pct[0], out["percentages"])
    assert 0 < sum(len(r) for r in rows[0]) < T * tc.n
    tr.generate_derivatives()
    assert out["cols_vs_batched_maxabs"] < 1e-9      # per-call drop-in vs the batched launch (another launch shape)
    tq = out["t_query"]
    A_t = tr.get(0, "A").reshape(T, 20, 20)[tq]
    assert np.allclose(np.array(out["A_t"]), A_t.ravel(), rtol=0, atol=1e-9)
    A_ref, _ = oracle.dynamics_derivs(om, ot, tr.get_snapshot(0, tq), list(range(tc.n)))
    assert np.linalg.norm(A_t.T - A_ref) <= 1e-4 * np.linalg.norm(A_ref)
    tr2 = capi.Traj(tc, T, 1)
    tr2.set_targets(0, target)
    tr2.set_keypoint_method(capi.KP_ITERATIVE_ERROR, 1, 50, jerk_thr=tc.jerk_thr, vel_thr=tc.vel_thr, iter_thr=1e-5)
    Uo, stats = tr2.optimise(snap0, U, 2, 1, lam=0.1)
    assert np.array_equal(Uo[0].ravel(), out["U_opt"])
    assert stats[0, 0] == out["initial_cost"] and stats[0, 1] == out["final_cost"] and int(stats[0, 2]) == out["iterations"]
    assert out["final_cost"] <= out["initial_cost"]
