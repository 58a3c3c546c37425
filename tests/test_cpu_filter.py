"""CPU: Optimiser::FilterDynamicsMatrices - the C++ oracle against an independent numpy
restatement, plus closed-form properties of both filters."""
import numpy as np
import pytest

import filter_numpy as fnp
from helpers import make, oracle_model, oracle_task
from oracle import oracle
from trajoptkp_b200 import capi


def _ilqr(T):
    tc = make("acrobot", kp_method=capi.KP_SET_INTERVAL, min_N=5)
    return tc, oracle.ILQR(oracle_model(tc), oracle_task(tc), T, threads=1)


@pytest.mark.parametrize("method,code", [("low_pass", 1), ("FIR", 2)])
def test_oracle_filter_matches_numpy_bit_exact(built, method, code):
    T = 37
    tc, il = _ilqr(T)
    rng = np.random.default_rng(5)
    A = rng.standard_normal(T * 16)
    il.set_A(A)
    il.set_filter(code)
    il.filter()
    want = fnp.filter_dynamics(A, tc.n, method)
    assert np.array_equal(il.get("A"), want)
    # position rows are never touched (Optimiser.cpp:345: i = dof .. 2 dof - 1)
    got = il.get("A").reshape(T, 4, 4)
    assert np.array_equal(got[:, :, :2], A.reshape(T, 4, 4)[:, :, :2])
    assert not np.array_equal(got[:, :, 2:], A.reshape(T, 4, 4)[:, :, 2:])


def test_custom_coefficients(built):
    T = 20
    tc, il = _ilqr(T)
    A = np.random.default_rng(6).standard_normal(T * 16)
    il.set_A(A)
    il.set_filter(2, fir=[0.25, 0.5, 0.25])
    il.filter()
    assert np.array_equal(il.get("A"), fnp.filter_dynamics(A, tc.n, "FIR", taps=(0.25, 0.5, 0.25)))
    il.set_A(A)
    il.set_filter(1, low_pass_a=0.6)
    il.filter()
    assert np.array_equal(il.get("A"), fnp.filter_dynamics(A, tc.n, "low_pass", a=0.6))


def test_filter_properties():
    # a constant signal passes the low-pass unchanged (unit DC gain) ...
    c = np.full(30, 3.5)
    assert np.allclose(fnp.low_pass(c), c, rtol=0, atol=1e-15)
    # ... and the default FIR once its 5 taps are filled (they sum to 1); before that the start is attenuated
    y = fnp.fir(c)
    assert np.allclose(y[4:], c[4:], rtol=0, atol=1e-15)
    assert np.allclose(y[:4], 3.5 * np.cumsum(fnp.DEFAULT_FIR)[:4])
    # linearity of both filters
    rng = np.random.default_rng(1)
    a, b = rng.standard_normal(25), rng.standard_normal(25)
    assert np.allclose(fnp.low_pass(2 * a - b), 2 * fnp.low_pass(a) - fnp.low_pass(b), atol=1e-12)
    assert np.allclose(fnp.fir(2 * a - b), 2 * fnp.fir(a) - fnp.fir(b), atol=1e-12)
    # impulse response of the FIR is its taps
    imp = np.zeros(10)
    imp[0] = 1.0
    assert np.array_equal(fnp.fir(imp)[:5], np.array(fnp.DEFAULT_FIR))


def test_generate_derivatives_applies_filter_last(built):
    """Optimiser.cpp:104-107: the filter runs after the cost derivatives, on the interpolated A."""
    T = 30
    tc, il = _ilqr(T)
    U = 0.1 * np.ones((T, tc.m))
    snap0 = tc.start_snapshot(start_pos=np.array([0.4, -0.1]))
    il.rollout(snap0, U)
    il.generate_derivatives()
    A_plain, lxx_plain = il.get("A"), il.get("l_xx")
    il.set_filter(1)
    il.generate_derivatives()
    assert np.array_equal(il.get("A"), fnp.filter_dynamics(A_plain, tc.n, "low_pass"))
    assert np.array_equal(il.get("l_xx"), lxx_plain)
