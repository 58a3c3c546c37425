#!/usr/bin/env python3
"""Where one world's step spends its cycles: clock64() at the stage boundaries of tk::step for world 0 of a
tokp_step_worlds launch (development build with -DTOKP_STAGE_CLOCKS in trajoptkp_b200/_build_prof; the product library
has none of this).

  python tools/stage_clocks.py --build                 # here: compile the instrumented library
  python tools/stage_clocks.py walker_run 3552 80 [warm-up steps] [control scale]   # on the GPU box: task, worlds, steps
"""
import ctypes as C
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROF = os.path.join(ROOT, "trajoptkp_b200", "_build_prof")
NAMES = {24: "kinematics: level 1", 25: "kinematics: level 2", 26: "kinematics: level 3", 27: "kinematics: levels 4+",
         0: "com_pos", 15: "crb", 1: "factor_ldl(M)", 2: "collision: compaction", 16: "collision: pair filter + primitives", 17: "collision: convex narrow phase", 3: "make_constraint",
         4: "com_vel + passive + bias", 5: "reference + actuation + acc_smooth", 6: "solver warm start",
         7: "update: forces, grad, cost", 8: "update: Hessian", 9: "update: Cholesky factor", 10: "update: solve",
         11: "line search", 12: "solver loop glue", 13: "euler"}


def build():
    """The product objects, with cuda_model.cu (k_step_worlds, the kernel this tool drives) recompiled with the clocks."""
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge
    ge.build()
    os.makedirs(PROF, exist_ok=True)
    objs = []
    for src in sorted(glob.glob(os.path.join(ge.CSRC, "*.cu")) + glob.glob(os.path.join(ge.CSRC, "*.cpp"))):
        base = os.path.basename(src)
        if base == "cuda_model.cu":
            obj = os.path.join(PROF, base + ".o")
            subprocess.check_call([ge.NVCC] + ge.NVCC_FLAGS + ["-DTOKP_STAGE_CLOCKS", "-c", src, "-o", obj])
        else:
            obj = os.path.join(ge.BUILD, base + ".o")
        objs.append(obj)
    subprocess.check_call([ge.NVCC] + ge.NVCC_FLAGS[:2] + ["-shared", "-o", os.path.join(PROF, "libtokp.so")] + objs + ["-lcudart"])


def main():
    if sys.argv[1] == "--build":
        return build()
    os.environ["TOKP_LIB"] = os.path.join(PROF, "libtokp.so")
    sys.path.insert(0, ROOT)
    import numpy as np
    from trajoptkp_b200 import capi, tasks
    task, count, nsteps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    tc = tasks.TaskConfig(task)
    m = tc.model
    rng = np.random.default_rng(0)
    snap = tc.start_snapshot()
    S = snap.size
    warm = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    cscale = float(sys.argv[5]) if len(sys.argv) > 5 else 0.3
    snaps = np.tile(snap, (count, 1))
    snaps[:, S - m.nu:] = cscale * rng.standard_normal((count, m.nu))
    L = capi.lib()

    def step_worlds(x, n):
        x = np.ascontiguousarray(x)
        out = np.zeros_like(x)
        st = np.zeros(count, dtype=np.int32)
        capi.check(L.tokp_step_worlds(m.h, capi._d(x), capi._d(out), count, n, -1, 0.0, capi._i(st)))
        return out

    snaps = step_worlds(snaps, warm)  # into a typical contact state
    snaps[:, S - m.nu:] = cscale * rng.standard_normal((count, m.nu))
    clk = (C.c_ulonglong * 32)()
    L.tokp_stage_clocks(clk, 1)
    step_worlds(snaps, nsteps)
    L.tokp_stage_clocks(clk, 0)
    steps = max(1, clk[22])
    tot = sum(clk[i] for i in NAMES)
    print("%s: %d worlds, %d steps of world 0; %.0f cycles per step; %.2f Newton updates, %.2f factorisations per step"
          % (task, count, steps, tot / steps, clk[20] / steps, clk[21] / steps))
    print("  narrow-phase pairs per step %.2f" % (clk[23] / steps))
    for i in [24, 25, 26, 27, 0, 15, 1, 16, 17, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13]:
        print("  %-36s %9.0f  %5.1f %%" % (NAMES[i], clk[i] / steps, 100.0 * clk[i] / tot))


if __name__ == "__main__":
    main()
