"""GPU: scheduling must not change results.  The same replans run (a) repeatedly, (b) with the warps of a block
walking the step independently instead of block-synchronously, (c) with other block sizes: identical arithmetic, so the
outputs must be identical bit for bit - a shared-memory race or a missing group barrier would show up here.  Other
lane counts change the summation order of the group reductions, so they are compared to tolerance."""
import os

import numpy as np
import pytest

from helpers import make, perturbed_start, random_controls, rel_fro
from trajoptkp_b200 import capi

pytestmark = pytest.mark.gpu


class _Env:
    def __init__(self, **kv):
        self.kv = {k: str(v) for k, v in kv.items()}

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in self.kv}
        os.environ.update(self.kv)

    def __exit__(self, *a):
        for k, v in self.old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def _replan(name, B, T, env):
    with _Env(**env):
        tc = make(name, kp_method=capi.KP_VELOCITY_CHANGE, min_N=1, max_N=10)
        snaps = np.stack([perturbed_start(tc, 40 + b, dq=0.05, dv=0.05) for b in range(B)])
        U = np.stack([random_controls(tc, T, 90 + b) for b in range(B)])
        tr = capi.Traj(tc, T, B)
        Uo, stats = tr.optimise(snaps, U, 2, 2, lam=0.1)
        out = {"U": Uo.copy(), "stats": stats[:, :4].copy()}
        for nm in ("A", "B", "X_old", "K", "k", "l_xx"):
            out[nm] = np.stack([tr.get(b, nm) for b in range(B)])
        return out


@pytest.mark.parametrize("name,B,T", [("walker_run", 24, 30), ("acrobot", 16, 40)])
def test_repeat_and_schedule_invariance(built, name, B, T):
    base = _replan(name, B, T, {})
    variants = [{}, {"TOKP_FD_BLOCKSYNC": 0, "TOKP_ROLL_BLOCKSYNC": 0}, {"TOKP_FD_BLOCKSYNC": 3, "TOKP_ROLL_BLOCKSYNC": 3},
                {"TOKP_WARPS": 1}, {"TOKP_WARPS": 4}, {"TOKP_SPREAD_WARPS": 1}]
    for env in variants:
        got = _replan(name, B, T, env)
        for k in base:
            assert np.array_equal(got[k], base[k]), (env, k)


def _derivs(lanes):
    with _Env(**({"TOKP_LANES": lanes} if lanes else {})):
        tc = make("walker_run", kp_method=capi.KP_SET_INTERVAL, min_N=5, max_N=5)
        B, T = 8, 30
        snaps = np.stack([perturbed_start(tc, 40 + b, dq=0.05, dv=0.05) for b in range(B)])
        U = np.stack([random_controls(tc, T, 90 + b) for b in range(B)])
        tr = capi.Traj(tc, T, B)
        cost = tr.rollout(snaps, U).copy()
        tr.generate_derivatives()
        return cost, np.stack([tr.get(b, "X_old") for b in range(B)]), np.stack([tr.get(b, "A") for b in range(B)])


@pytest.mark.parametrize("lanes", [8, 32])
def test_lane_count_changes_only_rounding(built, lanes):
    c0, X0, A0 = _derivs(None)
    c1, X1, A1 = _derivs(lanes)
    assert rel_fro(X1, X0) < 1e-9                      # per-step state tolerance of the north star
    assert np.allclose(c1, c0, rtol=1e-9)
    # FD Jacobians amplify rounding by 1/eps; ill-conditioned contact rows aside they agree
    per_row = [rel_fro(a, b) for a, b in zip(A1.reshape(-1, 18 * 18), A0.reshape(-1, 18 * 18))]
    assert np.median(per_row) < 1e-6
    assert np.mean(np.array(per_row) < 1e-4) > 0.9
