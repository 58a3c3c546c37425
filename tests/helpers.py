"""Shared helpers for the parity tests: build the oracle-side objects from a TaskConfig."""
import numpy as np

from oracle import oracle
from trajoptkp_b200 import tasks


def oracle_model(tc):
    return oracle.Model(tc.model.blob())


def oracle_task(tc, kp_method=None, min_N=None, max_N=None, target=None):
    return oracle.Task(tc.kind, tc.qpos_adr, tc.dof_adr, tc.ctrl_adr,
                       tc.target if target is None else target, tc.weight, tc.weight_terminal,
                       tc.kp_method if kp_method is None else kp_method,
                       tc.min_N if min_N is None else min_N, tc.max_N if max_N is None else max_N,
                       tc.jerk_thr, tc.vel_thr, tc.iter_thr, aux=tc.aux)


def random_controls(tc, T, seed, scale=0.3):
    rng = np.random.default_rng(seed)
    lim = tc.model.real("actuator_ctrlrange").reshape(-1, 2)[:tc.m]
    U = scale * rng.uniform(-1, 1, (T, tc.m)) * np.minimum(np.abs(lim[:, 0]), np.abs(lim[:, 1]))
    return U


def perturbed_start(tc, seed, dq=0.02, dv=0.01):
    """Synthetic initial-state perturbation: start positions += U(-dq,dq), velocities += N(0,dv)."""
    rng = np.random.default_rng(seed)
    start = tc.start_pos + rng.uniform(-dq, dq, tc.n)
    vel = dv * rng.standard_normal(tc.n)
    return tc.start_snapshot(start_pos=start, qvel=vel)


def rel_fro(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def make(name, **kw):
    return tasks.TaskConfig(name, **kw)
