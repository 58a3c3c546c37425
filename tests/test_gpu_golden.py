"""GPU: the CUDA path (through the C ABI) against the committed fixtures tests/golden/*.npz - the same seeded inputs
and oracle outputs the CPU suite pins.  Tolerances from BASELINE.json's north_star, written per assertion."""
import numpy as np
import pytest

from helpers import make, rel_fro
from test_cpu_golden import golden_rows, load
from trajoptkp_b200 import capi

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,ab_tol", [("acrobot", 1e-7), ("walker_run", 1e-4)])
def test_device_matches_golden(built, name, ab_tol):
    g = load(name)
    T = int(g["T"])
    tc = make(name, kp_method=int(g["method"]), min_N=int(g["min_N"]), max_N=int(g["max_N"]))
    tr = capi.Traj(tc, T, 2)   # two identical trajectories: batch slots must not interfere
    cost = tr.rollout(g["snap0"], g["U"])
    assert np.all(np.abs(cost - float(g["rollout_cost"])) <= 1e-9 * max(1.0, abs(float(g["rollout_cost"]))))
    X = g["X_old"]
    for b in range(2):
        assert np.abs(tr.get(b, "X_old") - X).max() / max(1.0, np.abs(X).max()) < 1e-9   # per-step states, 1e-9 rel
    rows, pct = tr.keypoints()
    assert rows[0] == golden_rows(g) and rows[1] == golden_rows(g)                      # key-points bit-exact
    assert np.array_equal(pct[0], g["pct"])
    tr.generate_derivatives()
    nx, m = tr.nx, tr.m
    A_ref, B_ref = g["A"].reshape(T, nx, nx), g["B"].reshape(T, m, nx)
    for b in range(2):
        A, B = tr.get(b, "A").reshape(T, nx, nx), tr.get(b, "B").reshape(T, m, nx)
        worst = max(max(rel_fro(A[t], A_ref[t]), rel_fro(B[t], B_ref[t])) for t in range(T))
        assert worst < ab_tol, worst                                                      # A/B relative Frobenius
    for f, tol in (("l_x", 1e-6), ("l_xx", 1e-6), ("l_u", 1e-6), ("l_uu", 1e-6)):
        assert rel_fro(tr.get(0, f), g[f]) < tol, f


@pytest.mark.parametrize("name", ["acrobot", "walker_run"])
def test_device_steps_match_golden(built, name):
    g = load(name)
    tc = make(name)
    m = tc.model
    sin = np.ascontiguousarray(g["steps_in"])
    out = np.zeros_like(sin)
    st = np.zeros(len(sin), dtype=np.int32)
    capi.check(capi.lib().tokp_step_worlds(m.h, capi._d(sin), capi._d(out), len(sin), int(g["nsteps"]), -1, 0.0, capi._i(st)))
    assert not st.any()
    ref = g["steps_out"]
    assert np.abs(out - ref).max() / max(1.0, np.abs(ref).max()) < 1e-9


@pytest.mark.parametrize("name", ["acrobot", "walker_run"])
def test_device_optimise_matches_golden_cost(built, name):
    g = load(name)
    T = int(g["T"])
    tc = make(name, kp_method=int(g["method"]), min_N=int(g["min_N"]), max_N=int(g["max_N"]))
    tr = capi.Traj(tc, T, 1)
    Uo, stats = tr.optimise(g["snap0"], g["U"], 3, 1, lam=0.1)
    assert abs(stats[0, 0] - float(g["initial_cost"])) <= 1e-9 * abs(float(g["initial_cost"]))
    assert abs(stats[0, 1] - float(g["final_cost"])) <= 0.01 * abs(float(g["final_cost"]))   # final cost within 1 %
