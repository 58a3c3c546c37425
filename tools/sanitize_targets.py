#!/usr/bin/env python3
"""Small launches for compute-sanitizer (racecheck / synccheck / memcheck): `fd_walker` = one batched FD-derivative
launch on a short walker trajectory (k_fd_plan, k_fd_worlds both phases, k_fd_bin, k_fd_dep_lanes, k_fd_assemble),
`rollout_reaching` = nominal + line-search rollouts of the reaching scene (k_rollout, generic-nv, convex collider),
`fd_box_sweep` = FD launch of the box_sweep scene (group kernels with stage sharing).
usage: compute-sanitizer --tool racecheck python tools/sanitize_targets.py fd_walker"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

from helpers import make, perturbed_start, random_controls  # noqa: E402
from trajoptkp_b200 import capi  # noqa: E402

what = sys.argv[1]
if what == "fd_walker":
    tc = make("walker_run", kp_method=capi.KP_SET_INTERVAL, min_N=2)
    B, T = 3, 8
    snaps = np.stack([perturbed_start(tc, 40 + b, dq=0.05, dv=0.05) for b in range(B)])
    U = np.stack([random_controls(tc, T, 90 + b) for b in range(B)])
    tr = capi.Traj(tc, T, B)
    tr.rollout(snaps, U)
    tr.generate_derivatives()
    print("fd_walker: worlds", tr.counters()[0], "|A| max", float(np.abs(tr.get(0, "A")).max()))
elif what == "fd_box_sweep":
    tc = make("box_sweep", kp_method=capi.KP_SET_INTERVAL, min_N=3)
    T = 6
    snap, goal = tc.init_snapshot(0)
    tr = capi.Traj(tc, T, 1)
    snap0, U, st = tr.init_controls_ee(snap[None], goal[None], pre_steps=1, setup_horizon=0)
    tr.rollout(snap0, U)
    tr.generate_derivatives()
    print("fd_box_sweep: worlds", tr.counters()[0], "|A| max", float(np.abs(tr.get(0, "A")).max()))
elif what == "rollout_reaching":
    tc = make("reaching", kp_method=capi.KP_SET_INTERVAL, min_N=2)
    B, T = 2, 5
    snaps = np.stack([tc.init_snapshot(b)[0] for b in range(B)])
    tr = capi.Traj(tc, T, B)
    U, st = tr.init_controls_gravcomp(snaps)
    Uo, stats = tr.optimise(snaps, U, 1, 1, lam=0.1)
    print("rollout_reaching: cost", stats[:, :2].tolist())
else:
    raise SystemExit("unknown target " + what)
