"""GPU: the two-pass constraint-row capacity scheme (fast pass with a small workspace, redo pass with maxefc for the
worlds that overflowed) must give the same results as a single full-capacity pass, for every world kernel."""
import os

import numpy as np
import pytest

from helpers import make, perturbed_start, random_controls
from trajoptkp_b200 import capi

pytestmark = pytest.mark.gpu


def _with_cap(cap, fn):
    old = os.environ.get("TOKP_EFC_CAP")
    os.environ["TOKP_EFC_CAP"] = str(cap)
    try:
        return fn()
    finally:
        if old is None:
            del os.environ["TOKP_EFC_CAP"]
        else:
            os.environ["TOKP_EFC_CAP"] = old


def test_step_worlds_redo_pass_matches_full_capacity(built):
    tc = make("walker_run")
    m = tc.model
    count = 32
    snaps = np.stack([perturbed_start(tc, 300 + i, dq=0.2, dv=0.5) for i in range(count)])
    snaps[:, 1] -= 0.15   # push the walker into the floor: many contacts

    def run():
        out = np.zeros_like(snaps)
        st = np.zeros(count, dtype=np.int32)
        capi.check(capi.lib().tokp_step_worlds(m.h, capi._d(snaps), capi._d(out), count, 12, -1, 0.0, capi._i(st)))
        return out, st

    full, st_full = _with_cap(10000, run)
    tiny, st_tiny = _with_cap(4, run)      # every contact world overflows the fast pass
    assert not (st_full & 8).any() and not (st_tiny & 8).any()
    assert np.array_equal(full, tiny)


def test_fd_and_rollout_redo_pass_match_full_capacity(built):
    tc = make("walker_run", kp_method=capi.KP_SET_INTERVAL, min_N=3)
    T = 24
    snap0 = perturbed_start(tc, 11)
    U = random_controls(tc, T, 5)

    def run():
        tr = capi.Traj(tc, T, 3)
        cost = tr.rollout(snap0, U)
        tr.keypoints()
        tr.generate_derivatives()
        ok, dJ = tr.backward_pass(0.1)
        costs = tr.forward_pass([(i / 6.0) ** 2 for i in range(1, 7)])
        return cost, tr.get(1, "A"), tr.get(2, "B"), tr.get(0, "X_old"), costs

    full = _with_cap(10000, run)
    tiny = _with_cap(4, run)
    for a, b in zip(full, tiny):
        assert np.array_equal(a, b)
