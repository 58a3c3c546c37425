"""GPU: tokp_step_worlds_static - the no-allocation variant of tokp_step_worlds (scratch in a static device array,
include/tokp.h) - against tokp_step_worlds on the same inputs, two calls in a row (the array keeps the last call's
contents), through the C ABI.  Collected last on purpose (file name).
Tolerance: per-step states 1e-9 relative (north_star); the two entry points run the same kernel, so the outputs are in
fact identical."""
import numpy as np
import pytest

from helpers import make
from trajoptkp_b200 import capi

pytestmark = pytest.mark.gpu

STATE_RTOL = 1e-9


def _inputs(tc, count, seed):
    m = tc.model
    rng = np.random.default_rng(seed)
    snaps = np.tile(tc.start_snapshot(), (count, 1))
    snaps[:, 1 + m.nq:1 + m.nq + m.nv] += 0.2 * rng.standard_normal((count, m.nv))
    snaps[:, m.S - m.nu:] = rng.uniform(-1, 1, (count, m.nu))
    return snaps


def _step(fn, m, snaps, nsteps):
    out = np.full_like(snaps, np.nan)
    st = np.full(len(snaps), -1, dtype=np.int32)
    capi.check(fn(m.h, capi._d(snaps), capi._d(out), len(snaps), nsteps, -1, 0.0, capi._i(st)))
    return out, st


def test_static_scratch_stepping_matches_allocating_entry_point(built):
    L = capi.lib()
    tc = make("walker_run")
    m = tc.model
    # two calls in a row with different sizes and inputs: the second must not see the first one's leftovers
    for count, nsteps, seed in ((96, 15, 1), (40, 25, 2)):
        snaps = _inputs(tc, count, seed)
        ref, st_ref = _step(L.tokp_step_worlds, m, snaps, nsteps)   # also uploads the model (the static variant does not)
        got, st = _step(L.tokp_step_worlds_static, m, snaps, nsteps)
        assert np.isfinite(got).all()
        assert np.array_equal(st, st_ref)
        err = np.abs(got - ref).max() / max(1.0, np.abs(ref).max())
        assert err <= STATE_RTOL, err
        assert np.abs(got - snaps).max() > 1e-3  # the worlds moved


def test_static_scratch_rejects_oversized_batches_and_cold_models(built):
    L = capi.lib()
    tc = make("walker_run")
    m = tc.model
    snaps = _inputs(tc, 4, 3)
    _step(L.tokp_step_worlds, m, snaps, 1)
    limit = (4 << 20) // (2 * 8 * m.S + 4)
    big = np.tile(snaps[:1], (limit + 1, 1))
    out = np.zeros_like(big)
    rc = L.tokp_step_worlds_static(m.h, capi._d(big), capi._d(out), len(big), 1, -1, 0.0, None)
    assert rc == 1 and b"static scratch holds" in L.tokp_last_error()  # TOKP_ERR_ARG
    ok = np.tile(snaps[:1], (limit, 1))
    out = np.zeros_like(ok)
    capi.check(L.tokp_step_worlds_static(m.h, capi._d(ok), capi._d(out), len(ok), 1, -1, 0.0, None))
    # a model that has never been on the device: the upload would allocate, so the call refuses
    cold = capi.Model(tc.model_path, tc.timestep)
    s = np.tile(tc.start_snapshot(), (2, 1))
    o = np.zeros_like(s)
    rc = L.tokp_step_worlds_static(cold.h, capi._d(s), capi._d(o), 2, 1, -1, 0.0, None)
    assert rc == 5 and b"not on the device" in L.tokp_last_error()  # TOKP_ERR_STATE
