"""GPU parity tests for key-points, interpolation, cost derivatives, the Riccati backward pass, the alpha-batched
forward pass, acceptance (with the reference's index quirks) and the full iLQR::Optimise loop."""
import numpy as np
import pytest

from helpers import make, oracle_model, oracle_task, perturbed_start, random_controls, rel_fro
from oracle import oracle
from trajoptkp_b200 import capi

pytestmark = pytest.mark.gpu

METHODS = [(capi.KP_SET_INTERVAL, 5, 20), (capi.KP_ADAPTIVE_JERK, 2, 10), (capi.KP_ADAPTIVE_ACCEL, 2, 10),
           (capi.KP_VELOCITY_CHANGE, 1, 20), (capi.KP_VELOCITY_CHANGE, 3, 8)]


def _setup(name, T, kp=None, min_N=None, max_N=None, seed=7, B=1):
    tc = make(name, kp_method=kp, min_N=min_N, max_N=max_N)
    snap0 = perturbed_start(tc, seed)
    U = random_controls(tc, T, seed + 1)
    tr = capi.Traj(tc, T, B)
    return tc, snap0, U, tr


@pytest.mark.parametrize("name", ["acrobot", "walker_run"])
@pytest.mark.parametrize("method,min_N,max_N", METHODS)
def test_keypoints_bit_exact(built, name, method, min_N, max_N):
    T = 50
    tc, snap0, U, tr = _setup(name, T, method, min_N, max_N)
    if method in (capi.KP_ADAPTIVE_JERK, capi.KP_ADAPTIVE_ACCEL):
        # thresholds that actually trigger on this trajectory
        thr = np.full(tc.n, 5.0 if method == capi.KP_ADAPTIVE_JERK else 0.002)
        tc.jerk_thr = thr
        tr.set_keypoint_method(method, min_N, max_N, jerk_thr=thr)
    if method == capi.KP_VELOCITY_CHANGE:
        thr = np.full(tc.n, 0.5)
        tc.vel_thr = thr
        tr.set_keypoint_method(method, min_N, max_N, vel_thr=thr)
    tr.rollout(snap0, U)
    rows, pct = tr.keypoints()
    X = tr.get(0, "X_old").reshape(T + 1, -1)
    ref_rows = oracle.keypoints(method, X, tc.n, T, tc.timestep, min_N, max_N, tc.jerk_thr, tc.vel_thr)
    assert rows[0] == ref_rows                                   # index lists, order and duplicates included
    assert (pct[0] == oracle.percentages(ref_rows, tc.n)).all()  # bit-exact percentages
    # structural invariants of the reference's own tests (src/tests/Keypoints_Test.cpp:10-33,81-114)
    assert rows[0][0] == list(range(tc.n))
    assert sorted(set(rows[0][T - 1])) == list(range(tc.n))
    if method != capi.KP_SET_INTERVAL:
        for j in range(tc.n):
            ts = [t for t in range(T) if j in rows[0][t]]
            assert max(b - a for a, b in zip(ts, ts[1:])) <= max_N


@pytest.mark.parametrize("name", ["acrobot", "walker_run"])
def test_interpolation_bit_exact(built, name):
    T = 40
    tc, snap0, U, tr = _setup(name, T, capi.KP_SET_INTERVAL, 3, 3)
    tr.rollout(snap0, U)
    rows, _ = tr.keypoints()
    tr.dynamics_derivs()
    nx, m = tr.nx, tr.m
    A0 = tr.get(0, "A").copy()
    B0 = tr.get(0, "B").copy()
    tr.interpolate()
    A1 = tr.get(0, "A").reshape(T, nx, nx)
    B1 = tr.get(0, "B")
    oracle.interpolate(rows[0], tc.n, m, A0, B0)
    assert (A1.ravel() == A0).all()
    assert (B1 == B0).all()
    # Keypoints_Test.cpp:272-289: A[1] == A[0] + (A[3]-A[0])/3 exactly (per element, same operation order)
    add = (A1[3] - A1[0]) / 3
    assert (A1[1] == A1[0] + 1.0 * add).all()


def test_iterative_error_keypoints(built):
    T = 33
    tc, snap0, U, tr = _setup("walker_run", T, capi.KP_ITERATIVE_ERROR, 2, 10)
    tc.iter_thr = 5e-3
    tr.set_keypoint_method(capi.KP_ITERATIVE_ERROR, 2, 10, iter_thr=5e-3)
    tr.rollout(snap0, U)
    rows, pct = tr.keypoints()
    ref = oracle.ILQR(oracle_model(tc), oracle_task(tc), T, threads=1)
    ref.rollout(snap0, U)
    ref.generate_keypoints()
    assert rows[0] == ref.keypoints()
    assert 0 < sum(len(r) for r in rows[0]) < T * tc.n


@pytest.mark.parametrize("name", ["acrobot", "walker_run"])
def test_cost_derivs_backward_forward(built, name):
    T = 40
    tc, snap0, U, tr = _setup(name, T, capi.KP_SET_INTERVAL, 2, 2)
    ref = oracle.ILQR(oracle_model(tc), oracle_task(tc), T, threads=4)
    c_ref = ref.rollout(snap0, U)
    ref.generate_derivatives()
    c_dev = tr.rollout(snap0, U)[0]
    tr.generate_derivatives()
    tol = 1e-4 if name == "walker_run" else 1e-7
    for nm in ("r_x", "r_u", "l_x", "l_xx", "l_u", "l_uu"):
        a, b = tr.get(0, nm), ref.get(nm)
        # l_*[T] is never written by the reference; compare t < T
        k = a.size // (T + 1)
        assert rel_fro(a[:T * k], b[:T * k]) < 1e-9, nm
    assert rel_fro(tr.get(0, "A"), ref.get("A")) < tol
    assert rel_fro(tr.get(0, "B"), ref.get("B")) < tol
    # backward pass from identical inputs: feed the oracle's derivatives to the device
    for nm in ("A", "B", "l_x", "l_xx", "l_u", "l_uu"):
        tr.set(0, nm, ref.get(nm))
    ref.set_lambda(0.1)
    ok_ref = ref.backward()
    ok, dJ = tr.backward_pass(0.1)
    assert bool(ok[0]) == ok_ref
    assert rel_fro(tr.get(0, "K"), ref.get("K")) < 1e-8
    assert rel_fro(tr.get(0, "k"), ref.get("k")) < 1e-8
    assert abs(dJ[0] - ref.scalars()["delta_J"]) <= 1e-8 * abs(ref.scalars()["delta_J"])
    # forward pass with identical gains
    tr.set(0, "K", ref.get("K"))
    tr.set(0, "k", ref.get("k"))
    alphas = [(i / 6.0) ** 2 for i in range(1, 7)]
    costs = tr.forward_pass(alphas)[0]
    ref_costs = np.array([ref.forward(i, a) for i, a in enumerate(alphas)])
    assert np.allclose(costs, ref_costs, rtol=1e-7, atol=1e-9), (costs, ref_costs)
    # acceptance: SaveBestRollout + UpdateNominal incl. the one-step shift and the zeroed first control
    best = int(np.argmin(ref_costs))
    ref.save_best(best, float(ref_costs[best]))
    tr.accept(0, best)
    for nm in ("X_old", "U_old", "resid"):
        a, b = tr.get(0, nm), ref.get(nm)
        assert np.allclose(a, b, rtol=1e-7, atol=1e-9), nm
    assert (tr.get(0, "U_old").reshape(T, -1)[0] == 0).all()          # quirk C.1-2
    sT = tr.get_snapshot(0, T)                                         # saved[T] <- never-written rollout slot:
    assert (sT[:tc.model.S - tc.model.nu] == 0).all()                  # state zeroed, its ctrl keeps the old value


@pytest.mark.parametrize("name,T,max_it", [("acrobot", 100, 6), ("walker_run", 40, 3)])
def test_optimise_final_cost(built, name, T, max_it):
    tc = make(name, kp_method=capi.KP_SET_INTERVAL, min_N=1)
    snap0 = perturbed_start(tc, 21)
    U = np.zeros((T, tc.m))
    ref = oracle.ILQR(oracle_model(tc), oracle_task(tc), T, threads=4)
    ref.set_lambda(0.1)
    U_ref = ref.optimise(snap0, U, max_it, 1)
    sc = ref.scalars()
    tr = capi.Traj(tc, T, 2)
    U_dev, stats = tr.optimise(snap0, U, max_it, 1, lam=0.1)
    for b in range(2):
        assert abs(stats[b, 0] - sc["initial_cost"]) <= 1e-9 * abs(sc["initial_cost"])
        assert abs(stats[b, 1] - sc["new_cost"]) <= 0.01 * abs(sc["new_cost"]), (stats[b], sc)   # final cost within 1 %
        assert int(stats[b, 2]) == int(sc["num_iterations"])
    assert stats[0, 1] < stats[0, 0]


@pytest.mark.parametrize("method", ["low_pass", "FIR"])
@pytest.mark.parametrize("name", ["acrobot", "walker_run"])
def test_filter_dynamics_bit_exact(built, name, method):
    """Optimiser::FilterDynamicsMatrices (a17): device kernel == oracle == numpy restatement, bit for bit, on the same A;
    batch entries are filtered independently."""
    import filter_numpy as fnp
    T = 33
    tc, snap0, U, _ = _setup(name, T, capi.KP_SET_INTERVAL, 2, 2)
    tr = capi.Traj(tc, T, 2)
    rng = np.random.default_rng(11)
    A = rng.standard_normal((2, T * (2 * tc.n) ** 2))
    for b in range(2):
        tr.set(b, "A", A[b])
    with pytest.raises(capi.TokpError):
        tr.filter_dynamics()                     # "Filtering method not recognised": no method set
    tr.set_filter(method)
    tr.filter_dynamics()
    ref = oracle.ILQR(oracle_model(tc), oracle_task(tc), T, threads=1)
    ref.set_filter(capi.Traj.FILTERS[method])
    for b in range(2):
        ref.set_A(A[b])
        ref.filter()
        got = tr.get(b, "A")
        assert np.array_equal(got, ref.get("A"))
        assert np.array_equal(got, fnp.filter_dynamics(A[b], tc.n, method))
    # custom coefficients
    tr.set(0, "A", A[0])
    tr.set_filter("FIR", fir=[0.3, 0.4, 0.3])
    tr.filter_dynamics()
    assert np.array_equal(tr.get(0, "A"), fnp.filter_dynamics(A[0], tc.n, "FIR", taps=(0.3, 0.4, 0.3)))


def test_generate_derivatives_with_filter(built):
    """GenerateDerivatives applies the filter after the cost derivatives (Optimiser.cpp:104-107)."""
    import filter_numpy as fnp
    T = 30
    tc, snap0, U, tr = _setup("acrobot", T, capi.KP_SET_INTERVAL, 2, 2)
    tr.rollout(snap0, U)
    tr.generate_derivatives()
    A_plain, lxx = tr.get(0, "A"), tr.get(0, "l_xx")
    tr.set_filter("low_pass", low_pass_a=0.25)
    tr.generate_derivatives()
    assert np.array_equal(tr.get(0, "A"), fnp.filter_dynamics(A_plain, tc.n, "low_pass"))
    assert np.array_equal(tr.get(0, "l_xx"), lxx)
    tr.set_filter("none")
    tr.generate_derivatives()
    assert np.array_equal(tr.get(0, "A"), A_plain)


@pytest.mark.parametrize("name", ["acrobot", "walker_run"])
def test_residual_derivs_per_call(built, name):
    """Differentiator::ResidualDerivatives drop-in (a2): one (b, t) per call == the batched phase == the oracle."""
    T = 20
    tc, snap0, U, tr = _setup(name, T, capi.KP_SET_INTERVAL, 2, 2)
    tr.rollout(snap0, U)
    tr.cost_derivs()
    rx_all = tr.get(0, "r_x").reshape(T + 1, 2 * tc.n, tc.nr)
    ru_all = tr.get(0, "r_u").reshape(T + 1, tc.m, tc.nr)
    om, ot = oracle_model(tc), oracle_task(tc)
    for t in (0, 7, T):
        rx, ru = tr.residual_derivs(0, t)
        assert np.array_equal(rx, rx_all[t]) and np.array_equal(ru, ru_all[t])
        rx_ref, ru_ref = oracle.residual_derivs(om, ot, tr.get_snapshot(0, t))
        assert rel_fro(rx.ravel(), np.asarray(rx_ref).T.ravel()) < 1e-9
        assert rel_fro(ru.ravel(), np.asarray(ru_ref).T.ravel()) < 1e-9
    # forward differences estimate the same derivative (identical for residuals linear in the state)
    rx_f, _ = tr.residual_derivs(0, 7, central=False)
    assert np.allclose(rx_f, rx_all[7], atol=1e-4)
    with pytest.raises(capi.TokpError):
        tr.residual_derivs(0, T + 1)


def test_dump_trajectory_writes_the_reference_csv_files(built, tmp_path):
    """FileHandler::SaveTrajecInformation (FileHandler.cpp:312-383) from a device trajectory."""
    from trajoptkp_b200 import sweep
    T = 30
    tc, snap0, U, tr = _setup("acrobot", T, capi.KP_SET_INTERVAL, 1, 1)
    tr.rollout(snap0, U)
    tr.generate_derivatives()
    root = sweep.dump_trajectory(tr, 0, str(tmp_path), "/acrobot_trial0")
    nx, m = tr.nx, tr.m
    A = np.array([[float(v) for v in line.rstrip(",\n").split(",")] for line in open(root + "/A_matrices.csv")])
    Bm = np.array([[float(v) for v in line.rstrip(",\n").split(",")] for line in open(root + "/B_matrices.csv")])
    X = np.array([[float(v) for v in line.rstrip(",\n").split(",")] for line in open(root + "/states.csv")])
    Uf = np.array([[float(v) for v in line.rstrip(",\n").split(",")] for line in open(root + "/controls.csv")])
    assert A.shape == (T - 1, nx * nx) and Bm.shape == (T - 1, nx * m) and X.shape == (T - 1, nx) and Uf.shape == (T - 1, m)
    # row-major A_t / B_t (the device stores them column-major), to the 6 printed digits
    A_dev = tr.get(0, "A").reshape(T, nx, nx).transpose(0, 2, 1)[:T - 1].reshape(T - 1, -1)
    B_dev = tr.get(0, "B").reshape(T, m, nx).transpose(0, 2, 1)[:T - 1].reshape(T - 1, -1)
    assert np.allclose(A, A_dev, rtol=1e-5, atol=1e-9) and np.allclose(Bm, B_dev, rtol=1e-5, atol=1e-9)
    assert np.allclose(X, tr.get(0, "X_old").reshape(T + 1, nx)[:T - 1], rtol=1e-5, atol=1e-6)
    assert np.allclose(Uf, np.asarray(U).reshape(T, m)[:T - 1], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("method,min_N,max_N", [(capi.KP_VELOCITY_CHANGE, 2, 20), (capi.KP_SET_INTERVAL, 5, 5)])
def test_device_keypoint_statistics_equal_the_host_readback(built, method, min_N, max_N):
    """k_kp_stats (the derivative phase no longer copies the key-point masks to the host): the average percentage of
    derivatives in tokp_optimise's stats and the FD world / Jacobian counters equal what the host computes from
    tokp_keypoints - including velocity_change's last row (every dof appended again, KeyPointGenerator.cpp:724-727)."""
    T, B = 40, 3
    tc = make("walker_run", kp_method=method, min_N=min_N, max_N=max_N)
    if method == capi.KP_VELOCITY_CHANGE:
        tc.vel_thr = np.full(tc.n, 0.5)
    snaps = np.ascontiguousarray(np.stack([perturbed_start(tc, 30 + b) for b in range(B)]))
    U = np.ascontiguousarray(np.stack([np.asarray(random_controls(tc, T, 40 + b)).reshape(T, tc.m) for b in range(B)]))
    tr = capi.Traj(tc, T, B)
    Uo, stats = np.zeros((B, T, tc.m)), np.zeros((B, 8))
    capi.check(capi.lib().tokp_optimise(tr.h, capi._d(snaps), capi._d(U), 1, 1, None, capi._d(Uo), capi._d(stats)))
    counters = tr.counters()
    # the key-points of that iteration: the same generator on the same nominal trajectory, read back on the host
    tr2 = capi.Traj(tc, T, B)
    capi.check(capi.lib().tokp_rollout(tr2.h, capi._d(snaps), capi._d(U), None))
    rows, pct = tr2.keypoints()
    worlds = cols = 0
    for b in range(B):
        assert stats[b, 4] == pytest.approx(np.asarray(pct[b]).sum() / tc.n, rel=1e-14), (b, stats[b, 4], pct[b])
        for row in rows[b]:
            c = set(row)
            if c:
                worlds += 1 + 2 * len([j for j in c if j < tc.m]) + 4 * len(c)
                cols += len(c)
    assert counters[0] == worlds
    assert counters[1] == int(cols / tc.n * 1e6)
