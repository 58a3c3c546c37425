#!/usr/bin/env python3
"""Export (snapshot, next state, A, B) triples from a REAL MuJoCo install, to pin the oracle at the MuJoCo boundary.

Unrunnable in the build container (no `mujoco` wheel, no network) - ships so that anyone with MuJoCo >= 3.1 can close
the "parity unpinned" gap recorded in DESIGN.md section 2:

    python tools/export_mujoco_snapshots.py assets/models/walker_plane.xml 0.005 out.npz
    # then: TOKP_MUJOCO_PINS=out.npz python -m pytest tests/test_cpu_oracle_physics.py

Snapshot layout = MuJoCoHelper::CpMjData (src/PhysicsSimulators/MuJoCoHelper.cpp:732-741):
[time | qpos | qvel | qacc | qacc_warmstart | qfrc_applied | xfrc_applied | ctrl]."""
import sys

import numpy as np


def snapshot(m, d):
    return np.concatenate([[d.time], d.qpos, d.qvel, d.qacc, d.qacc_warmstart, d.qfrc_applied, d.xfrc_applied.ravel(), d.ctrl])


def main():
    import mujoco  # noqa: deliberately imported here
    xml, dt, out = sys.argv[1], float(sys.argv[2]), sys.argv[3]
    m = mujoco.MjModel.from_xml_path(xml)
    m.opt.timestep = dt
    d = mujoco.MjData(m)
    rng = np.random.default_rng(0)
    snaps, nexts, As, Bs = [], [], [], []
    for k in range(16):
        mujoco.mj_resetData(m, d)
        d.qvel[:] = 0.3 * rng.standard_normal(m.nv)
        d.ctrl[:] = rng.uniform(-0.5, 0.5, m.nu)
        for _ in range(20 + 10 * k):
            mujoco.mj_step(m, d)
        snaps.append(snapshot(m, d))
        A = np.zeros((2 * m.nv, 2 * m.nv))
        B = np.zeros((2 * m.nv, m.nu))
        it, tol = m.opt.iterations, m.opt.tolerance
        m.opt.iterations, m.opt.tolerance = 5, 0.0   # MuJoCoHelper::InitModelForFiniteDifferencing (:925-931)
        mujoco.mjd_transitionFD(m, d, 1e-6, True, A, B, None, None)
        m.opt.iterations, m.opt.tolerance = it, tol
        As.append(A)
        Bs.append(B)
        mujoco.mj_step(m, d)
        nexts.append(snapshot(m, d))
    np.savez_compressed(out, snap=np.array(snaps), next=np.array(nexts), A=np.array(As), B=np.array(Bs),
                        mujoco_version=mujoco.__version__)


if __name__ == "__main__":
    main()
