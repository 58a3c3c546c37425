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
