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


def _ab(name, B, T, share, kp=capi.KP_SET_INTERVAL, min_N=3, lanes=0):
    tc = make(name, kp_method=kp, min_N=min_N, max_N=10)
    if tc.kind in (capi.TASK_BOXSWEEP, capi.TASK_PUSH):
        snaps, goals = zip(*[tc.init_snapshot(b) for b in range(B)])
        tr = capi.Traj(tc, T, B)
        snap0, U, st = tr.init_controls_ee(np.stack(snaps), np.stack(goals), pre_steps=1, setup_horizon=0)
    else:
        snap0 = np.stack([perturbed_start(tc, 40 + b, dq=0.05, dv=0.05) for b in range(B)])
        U = np.stack([random_controls(tc, T, 90 + b) for b in range(B)])
        tr = capi.Traj(tc, T, B)
    tr.set_option("fd_stage_sharing", share)
    tr.set_option("fd_lanes", lanes)
    tr.rollout(snap0, U)
    tr.generate_derivatives()
    return (np.stack([tr.get(b, "A") for b in range(B)]), np.stack([tr.get(b, "B") for b in range(B)]),
            tr.counters()[0])


@pytest.mark.parametrize("name,B,T,kp", [("walker_run", 12, 40, capi.KP_SET_INTERVAL), ("acrobot", 8, 50, capi.KP_SET_INTERVAL),
                                         ("box_sweep", 3, 150, capi.KP_SET_INTERVAL), ("reaching", 4, 30, capi.KP_SET_INTERVAL),
                                         ("walker_run", 6, 33, capi.KP_ITERATIVE_ERROR)])
def test_stage_sharing_changes_no_bit(built, name, B, T, kp):
    """the control / velocity worlds of a key-point take the nominal world's position (and velocity) stage
    and its factor of M + h D, as mj_stepSkip does in the reference (Differentiator.cpp:118,155,253,276).  The shared
    values are the ones each world would have computed itself, so A and B must not change in a single bit."""
    A1, B1, w1 = _ab(name, B, T, 1, kp)
    A0, B0, w0 = _ab(name, B, T, 0, kp)
    assert w1 == w0 and w1 > 0
    assert np.array_equal(A1, A0)
    assert np.array_equal(B1, B0)
    assert np.abs(A1).max() > 0


@pytest.mark.parametrize("lanes", [1, 2])
@pytest.mark.parametrize("name,B,T", [("walker_run", 12, 40), ("acrobot", 8, 50)])
def test_lane_per_world_kernel_changes_only_rounding(built, name, B, T, lanes):
    """The worlds of small hinge / slide models run one world per lane (k_fd_dep_lanes; fd_lanes 1: the control / velocity
    worlds, 2: every world after a position-stage-only group kernel): sequential sums instead of butterfly reductions, so
    A / B agree with the group kernels up to rounding amplified by 1 / eps."""
    A1, B1, w1 = _ab(name, B, T, 1, lanes=lanes)
    A0, B0, w0 = _ab(name, B, T, 1, lanes=0)
    assert w1 == w0 and w1 > 0
    nx = A1.shape[-1] if A1.ndim == 3 else None
    per_row = [rel_fro(a, b) for a, b in zip(A1.reshape(B * T, -1), A0.reshape(B * T, -1)) if np.abs(b).max() > 0]
    assert np.median(per_row) < 1e-6, np.median(per_row)
    assert np.mean(np.array(per_row) < 1e-4) > 0.9
    per_row_b = [rel_fro(a, b) for a, b in zip(B1.reshape(B * T, -1), B0.reshape(B * T, -1)) if np.abs(b).max() > 0]
    assert np.median(per_row_b) < 1e-6, np.median(per_row_b)
