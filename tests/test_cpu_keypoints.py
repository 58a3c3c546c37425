"""CPU tests of the oracle's key-point rules and interpolation against (a) the known answers the reference's own
tests / published data hold (SURVEY 8(c) pins 1-4), and (b) an independent pure-Python restatement (kp_numpy.py).
Integer outputs must be bit-exact."""
import numpy as np
import pytest

import kp_numpy
from oracle import oracle


def _traj(T, n, seed, smooth=True):
    rng = np.random.default_rng(seed)
    t = np.linspace(0, 1, T + 1)[:, None]
    X = np.zeros((T + 1, 2 * n))
    for j in range(n):
        f = rng.uniform(0.5, 6.0)
        X[:, j] = np.sin(2 * np.pi * f * t[:, 0] + rng.uniform(0, 6))
        X[:, n + j] = rng.uniform(0.2, 3.0) * np.cos(2 * np.pi * f * t[:, 0]) + (0 if smooth else 0.3 * rng.standard_normal(T + 1))
    return X


# ---- pin (1): set_interval rows, reference Keypoints_Test.cpp:81-114 ------------------------------------------
@pytest.mark.parametrize("T,n,min_N", [(10, 2, 2), (100, 9, 5), (500, 2, 5), (20, 9, 20), (7, 3, 1)])
def test_set_interval_rows(T, n, min_N):
    X = np.zeros((T + 1, 2 * n))
    rows = oracle.keypoints(oracle.KP_SET_INTERVAL, X, n, T, 0.01, min_N, 50, np.ones(n), np.ones(n))
    assert len(rows) == T
    full = list(range(n))
    for t in range(T - 1):
        assert rows[t] == (full if t % min_N == 0 else [])
    assert rows[T - 1] == full
    assert rows == kp_numpy.set_interval(T, n, min_N)


# ---- pin (4): published %-derivative figures for set_interval (PaperData summaries) -------------------
@pytest.mark.parametrize("T,min_N,pct", [(500, 5, 20.2), (1500, 5, 300 / 1500 * 100 + 1 / 1500 * 100), (1000, 5, 20.1),
                                         (500, 1000, 0.4), (1000, 1000, 0.2), (20, 5, 25.0), (20, 20, 10.0),
                                         (40, 20, 7.5)])
def test_set_interval_percentages_known_answers(T, min_N, pct):
    n = 3
    rows = oracle.keypoints(oracle.KP_SET_INTERVAL, np.zeros((T + 1, 2 * n)), n, T, 0.01, min_N, 50, np.ones(n), np.ones(n))
    p = oracle.percentages(rows, n)
    assert np.allclose(p, pct, rtol=0, atol=1e-9), (p, pct)
    assert abs(round(1500 * 20.0667 / 100) - 301) <= 0  # 20.0667 % of 1500 == 301 rows (sanity on the table value)


# ---- pin (2): adaptive methods: first and last rows full, per-dof gap <= max_N (Keypoints_Test.cpp:10-33) --------
def _check_structure(rows, n, T, max_N, allow_dup_last=False):
    full = list(range(n))
    assert rows[0] == full
    last = sorted(set(rows[T - 1])) if allow_dup_last else rows[T - 1]
    assert last == full
    for j in range(n):
        ts = [t for t in range(T) if j in rows[t]]
        gaps = np.diff(ts)
        assert gaps.max() <= max_N, (j, gaps.max())


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
@pytest.mark.parametrize("min_N,max_N", [(1, 10), (2, 7), (5, 50)])
def test_adaptive_jerk_matches_independent_restatement(seed, min_N, max_N):
    T, n, dt = 120, 4, 0.004
    X = _traj(T, n, seed, smooth=False)
    thr = np.array([2e3, 1e4, 5e4, 1e6])
    rows = oracle.keypoints(oracle.KP_ADAPTIVE_JERK, X, n, T, dt, min_N, max_N, thr, thr)
    assert rows == kp_numpy.adaptive_jerk(X, n, T, dt, min_N, max_N, thr)
    _check_structure(rows, n, T, max_N)


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
@pytest.mark.parametrize("min_N,max_N", [(1, 20), (3, 9), (5, 100)])
def test_velocity_change_matches_independent_restatement(seed, min_N, max_N):
    T, n = 150, 5
    X = _traj(T, n, seed, smooth=(seed % 2 == 0))
    thr = np.array([0.5, 2.0, 6.0, 10.0, 1e9])
    rows = oracle.keypoints(oracle.KP_VELOCITY_CHANGE, X, n, T, 0.005, min_N, max_N, thr, thr)
    ref = kp_numpy.velocity_change(X, n, T, min_N, max_N, thr)
    assert rows == ref
    _check_structure(rows, n, T, max_N, allow_dup_last=True)
    # all dofs appended again to row T-1 (duplicates inflate the percentages, KeyPointGenerator.cpp:724-727,814-838)
    assert rows[T - 1][-n:] == list(range(n))
    assert np.array_equal(oracle.percentages(rows, n), kp_numpy.percentages(ref, n))


def test_velocity_change_turning_point_rule_is_dead_for_min_N_1():
    """last_vel_direction is only updated while cnt < min_N, so with min_N = 1 the turning-point rule never fires
    (KeyPointGenerator.cpp:684-697): a sign flip in the velocity slope alone must not create a key-point."""
    T, n = 40, 1
    X = np.zeros((T + 1, 2))
    X[:, 1] = 1e-3 * np.abs(np.arange(T + 1) - 20)  # V-shaped velocity, tiny magnitude: no threshold crossing
    rows = oracle.keypoints(oracle.KP_VELOCITY_CHANGE, X, n, T, 0.005, 1, 1000, np.array([1e9]), np.array([1e9]))
    assert [t for t in range(T) if rows[t]] == [0, T - 1]


# ---- pin (3): interpolation identity, Keypoints_Test.cpp:272-289 ------------------------------------------------
def test_interpolation_identity_bit_exact():
    n, m, T = 3, 2, 7
    nx = 2 * n
    rng = np.random.default_rng(5)
    A = rng.standard_normal((T, nx, nx))
    B = rng.standard_normal((T, m, nx))
    rows = [[0, 1, 2], [], [], [0, 1, 2], [1], [], [0, 1, 2]]
    A0, B0 = A.copy(), B.copy()
    oracle.interpolate(rows, n, m, A, B)
    # the reference test's expression: A[1] == A[0] + (A[3]-A[0])/3 ; A[2] == A[0] + 2*((A[3]-A[0])/3)
    for k in (1, 2):
        assert np.array_equal(A[k], A0[0] + k * ((A0[3] - A0[0]) / 3))
        assert np.array_equal(B[k], B0[0] + k * ((B0[3] - B0[0]) / 3))
    # dof 1 has a key-point at t=4; dofs 0 and 2 interpolate 3 -> 6
    for c in (0, 2, 3, 5):
        for k in (4, 5):
            assert np.array_equal(A[k, c], A0[3, c] + (k - 3) * ((A0[6, c] - A0[3, c]) / 3))
    for c in (1, 4):
        assert np.array_equal(A[4, c], A0[4, c])
        assert np.array_equal(A[5, c], A0[4, c] + 1 * ((A0[6, c] - A0[4, c]) / 2))
    # B column keyed by DoF index i < m (KeyPointGenerator.cpp:927-931): column 1 follows dof 1's key-points
    assert np.array_equal(B[5, 1], B0[4, 1] + 1 * ((B0[6, 1] - B0[4, 1]) / 2))
    # key-point rows untouched
    for t in (0, 3, 6):
        assert np.array_equal(A[t], A0[t]) and np.array_equal(B[t], B0[t])


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_interpolation_matches_independent_restatement(seed):
    n, m, T = 5, 3, 60
    nx = 2 * n
    rng = np.random.default_rng(seed)
    X = _traj(T, n, seed, smooth=False)
    thr = np.array([0.5, 2.0, 6.0, 10.0, 3.0])
    rows = oracle.keypoints(oracle.KP_VELOCITY_CHANGE, X, n, T, 0.005, 2, 11, thr, thr)
    A = rng.standard_normal((T, nx, nx))
    B = rng.standard_normal((T, m, nx))
    A2, B2 = A.copy(), B.copy()
    oracle.interpolate(rows, n, m, A, B)
    kp_numpy.interpolate(rows, n, m, A2, B2)
    assert np.array_equal(A, A2)
    assert np.array_equal(B, B2)
