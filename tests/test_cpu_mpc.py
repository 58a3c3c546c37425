"""CPU: host logic of the synchronous MPC harness (GenTestingData.cpp:422-751) - control-buffer shift,
best-matching-state search, residual/cost restatement against the oracle's task code, result files."""
import numpy as np
import pytest

from helpers import make, oracle_model, oracle_task
from oracle import oracle
from trajoptkp_b200 import capi, mpc


def test_shift_controls_repeats_last():
    U = np.arange(12.0).reshape(6, 2)
    assert np.array_equal(mpc.shift_controls(U, 0), U)
    S = mpc.shift_controls(U, 2)
    assert S.shape == U.shape and np.array_equal(S[:4], U[2:]) and np.array_equal(S[4], U[5]) and np.array_equal(S[5], U[5])


def test_best_matching_state_index_rules():
    H, nx = 6, 4
    X = np.arange((H + 1) * nx, dtype=float).reshape(H + 1, nx)
    assert mpc.best_matching_state_index(X, X[3], H) == 3
    # first minimum wins (strict <), the last searched index is horizon - 2
    X2 = X.copy()
    X2[4] = X2[2]
    assert mpc.best_matching_state_index(X2, X2[2], H) == 2
    assert mpc.best_matching_state_index(X, X[5] + 0.4, H) == 4
    # nothing within 1000: the default (clipped to horizon - 1)
    assert mpc.best_matching_state_index(X + 1e6, X[0], H, default_index=9) == H - 1


@pytest.mark.parametrize("name", ["acrobot", "walker_run"])
def test_host_residuals_match_oracle_cost(built, name):
    """Final-cost replay uses the host restatement of MT::Residuals / CostFunction: same cost as the oracle's rollout
    accounting (cost after every step, terminal weights on the last)."""
    tc = make(name)
    om, ot = oracle_model(tc), oracle_task(tc)
    T = 12
    rng = np.random.default_rng(3)
    U = 0.2 * rng.standard_normal((T, tc.m))
    snap0 = tc.start_snapshot()
    il = oracle.ILQR(om, ot, T, threads=1)
    want = il.rollout(snap0, U)
    m = tc.model
    d = oracle.Data(om)
    d.load(snap0)
    cost = 0.0
    for t in range(T):
        s = d.save()
        s[m.S - m.nu:][tc.ctrl_adr] = U[t]
        d.load(s)
        d.step()
        s = d.save()
        res = mpc.task_residuals(tc, s[1:1 + m.nq], s[1 + m.nq:1 + m.nq + m.nv], s[m.S - m.nu:], tc.target)
        cost += mpc.task_cost(tc, res, t == T - 1)
    assert abs(cost - want) <= 1e-12 * max(1.0, abs(want))


def test_mpc_result_files(built, tmp_path):
    import time

    import pandas as pd
    import yaml
    tc = make("walker_run", kp_method=capi.KP_ADAPTIVE_JERK, min_N=2, max_N=20)
    rng = np.random.default_rng(0)
    res = {"summary": np.abs(rng.standard_normal((3, 9))), "per_replan": np.abs(rng.standard_normal((5, 3, 5))),
           "trials": [0, 1, 2]}
    when = time.struct_time((2024, 9, 17, 21, 33, 0, 1, 261, -1))
    d = mpc.write_mpc_results(str(tmp_path), tc, 20, res, when=when)
    assert d.endswith("TestingData/iLQR/walker_run_synchronus_mpc_20240917_2133")
    df = pd.read_csv(d + "/summary.csv")
    assert list(df.columns) == ["Final cost", "Final dist", "Average dofs", "Average optimisation time (ms)",
                                "Average percent derivs", "Average time derivs (ms)", "Average time BP (ms)",
                                "Average time FP (ms)", "Average surprise"]
    assert len(df) == 3
    t1 = pd.read_csv(d + "/1.csv")
    assert list(t1.columns) == ["Optimisation time (ms)", "num dofs", "% derivs", "derivs time (ms)", "BP time (ms)",
                                "FP time (ms)", "surprise", "expected", "new cost"] and len(t1) == 5
    y = yaml.safe_load(open(d + "/summary.yaml"))
    assert y["keypoint_name"] == "AJ_2_20" and y["controls_noise"] == 0.5 and len(y["jerk_thresholds"]) == tc.n
