"""ORACLE (test infrastructure only): ctypes binding of oracle/_build/liboracle.so.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")

KP_SET_INTERVAL, KP_ADAPTIVE_JERK, KP_ADAPTIVE_ACCEL, KP_ITERATIVE_ERROR, KP_VELOCITY_CHANGE = range(5)
STAGE_NONE, STAGE_POS, STAGE_VEL = 0, 1, 2


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".h", ".cpp"))]
    if (not force and os.path.exists(_LIB_PATH)
            and all(os.path.getmtime(_LIB_PATH) >= os.path.getmtime(s) for s in srcs)):
        return _LIB_PATH
    subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


class TaskDesc(C.Structure):
    """Mirror of tokp_task_desc (include/tokp.h) / ok::TaskDesc."""
    _fields_ = [("kind", C.c_int), ("n", C.c_int), ("m", C.c_int),
                ("qpos_adr", C.POINTER(C.c_int)), ("dof_adr", C.POINTER(C.c_int)), ("ctrl_adr", C.POINTER(C.c_int)),
                ("num_resid", C.c_int),
                ("resid_target", C.POINTER(C.c_double)), ("resid_weight", C.POINTER(C.c_double)),
                ("resid_weight_terminal", C.POINTER(C.c_double)),
                ("kp_method", C.c_int), ("min_N", C.c_int), ("max_N", C.c_int),
                ("jerk_thresholds", C.POINTER(C.c_double)), ("velchange_thresholds", C.POINTER(C.c_double)),
                ("iterative_error_threshold", C.c_double),
                ("num_target", C.c_int), ("num_aux", C.c_int), ("aux", C.POINTER(C.c_int))]


_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def _d(a):
    return a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        L.ok_last_error.restype = C.c_char_p
        L.ok_model_create.restype = C.c_void_p
        L.ok_model_create.argtypes = [C.c_void_p, C.c_int64]
        L.ok_model_destroy.argtypes = [C.c_void_p]
        L.ok_snapshot_size.argtypes = [C.c_void_p]
        L.ok_data_create.restype = C.c_void_p
        L.ok_data_create.argtypes = [C.c_void_p]
        L.ok_data_destroy.argtypes = [C.c_void_p]
        L.ok_data_load_snapshot.argtypes = [C.c_void_p, _dp]
        L.ok_data_save_snapshot.argtypes = [C.c_void_p, _dp]
        L.ok_data_step.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double]
        L.ok_data_forward.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double]
        L.ok_data_get.argtypes = [C.c_void_p, C.c_char_p, _dp, C.c_int]
        L.ok_integrate_pos.argtypes = [C.c_void_p, _dp, _dp, C.c_double]
        L.ok_differentiate_pos.argtypes = [C.c_void_p, _dp, C.c_double, _dp, _dp]
        L.ok_flops_step.restype = C.c_uint64
        L.ok_flops_step.argtypes = [C.c_void_p, _dp, C.c_int, C.c_int, C.c_double]
        L.ok_dynamics_derivs.argtypes = [C.c_void_p, C.POINTER(TaskDesc), _dp, _ip, C.c_int, C.c_int, C.c_double,
                                         C.c_int, C.c_double, _dp, _dp]
        L.ok_flops_dynamics_derivs.restype = C.c_uint64
        L.ok_flops_dynamics_derivs.argtypes = [C.c_void_p, C.POINTER(TaskDesc), _dp, _ip, C.c_int, C.c_int, C.c_double]
        L.ok_residual_derivs.argtypes = [C.c_void_p, C.POINTER(TaskDesc), _dp, C.c_int, C.c_double, _dp, _dp]
        L.ok_residuals.argtypes = [C.c_void_p, C.POINTER(TaskDesc), _dp, _dp]
        L.ok_keypoints.argtypes = [C.c_int, _dp, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, _dp, _dp, _ip, _ip, C.c_int]
        L.ok_interpolate.argtypes = [_ip, _ip, C.c_int, C.c_int, C.c_int, _dp, _dp]
        L.ok_percentages.argtypes = [_ip, _ip, C.c_int, C.c_int, _dp]
        L.ok_init_sequence_ee.argtypes = [C.c_void_p, C.POINTER(TaskDesc), _dp, _dp, C.c_int, C.c_int, _dp, _dp, _dp]
        L.ok_ilqr_create.restype = C.c_void_p
        L.ok_ilqr_create.argtypes = [C.c_void_p, C.POINTER(TaskDesc), C.c_int, C.c_int]
        L.ok_ilqr_destroy.argtypes = [C.c_void_p]
        L.ok_ilqr_set_lambda.argtypes = [C.c_void_p, C.c_double]
        L.ok_ilqr_set_fd_model_opts.argtypes = [C.c_void_p, C.c_int]
        L.ok_ilqr_set_filter.argtypes = [C.c_void_p, C.c_int, C.c_double, _dp, C.c_int]
        L.ok_ilqr_filter.argtypes = [C.c_void_p]
        L.ok_ilqr_set_A.argtypes = [C.c_void_p, _dp]
        L.ok_ilqr_rollout.restype = C.c_double
        L.ok_ilqr_rollout.argtypes = [C.c_void_p, _dp, _dp]
        for f in ("ok_ilqr_generate_keypoints", "ok_ilqr_dynamics_derivs", "ok_ilqr_generate_derivatives"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.ok_ilqr_backward.argtypes = [C.c_void_p]
        L.ok_ilqr_forward.restype = C.c_double
        L.ok_ilqr_forward.argtypes = [C.c_void_p, C.c_int, C.c_double]
        L.ok_ilqr_save_best.argtypes = [C.c_void_p, C.c_int, C.c_double]
        L.ok_ilqr_iteration.argtypes = [C.c_void_p, _ip]
        L.ok_ilqr_optimise.argtypes = [C.c_void_p, _dp, _dp, C.c_int, C.c_int, _dp]
        L.ok_ilqr_get.argtypes = [C.c_void_p, C.c_char_p, _dp, C.c_int]
        L.ok_ilqr_get_keypoints.argtypes = [C.c_void_p, _ip, _ip, C.c_int]
        L.ok_ilqr_set_nominal.argtypes = [C.c_void_p, _dp, _dp, _dp]
        _lib = L
    return _lib


class Task:
    """Owns the numpy buffers behind a TaskDesc."""

    def __init__(self, kind, qpos_adr, dof_adr, ctrl_adr, target, weight, weight_terminal, kp_method, min_N, max_N,
                 jerk_thr, vel_thr, iter_thr, aux=None):
        self.qpos_adr = np.ascontiguousarray(qpos_adr, dtype=np.int32)
        self.dof_adr = np.ascontiguousarray(dof_adr, dtype=np.int32)
        self.ctrl_adr = np.ascontiguousarray(ctrl_adr, dtype=np.int32)
        self.target = np.ascontiguousarray(target, dtype=np.float64)
        self.weight = np.ascontiguousarray(weight, dtype=np.float64)
        self.weight_terminal = np.ascontiguousarray(weight_terminal, dtype=np.float64)
        self.jerk_thr = np.ascontiguousarray(jerk_thr, dtype=np.float64)
        self.vel_thr = np.ascontiguousarray(vel_thr, dtype=np.float64)
        self.kind, self.kp_method, self.min_N, self.max_N, self.iter_thr = kind, kp_method, min_N, max_N, iter_thr
        self.aux = np.ascontiguousarray([] if aux is None else aux, dtype=np.int32)
        self.n, self.m, self.nr = len(self.qpos_adr), len(self.ctrl_adr), len(self.weight)

    def desc(self):
        d = TaskDesc()
        d.kind, d.n, d.m = self.kind, self.n, self.m
        d.qpos_adr, d.dof_adr, d.ctrl_adr = _i(self.qpos_adr), _i(self.dof_adr), _i(self.ctrl_adr)
        d.num_resid = self.nr
        d.resid_target, d.resid_weight, d.resid_weight_terminal = _d(self.target), _d(self.weight), _d(self.weight_terminal)
        d.kp_method, d.min_N, d.max_N = self.kp_method, self.min_N, self.max_N
        d.jerk_thresholds, d.velchange_thresholds = _d(self.jerk_thr), _d(self.vel_thr)
        d.iterative_error_threshold = self.iter_thr
        d.num_target = len(self.target) if len(self.target) != self.nr else 0
        d.num_aux = len(self.aux)
        d.aux = _i(self.aux) if len(self.aux) else None
        return d


class Model:
    def __init__(self, blob: bytes):
        self._blob = np.frombuffer(blob, dtype=np.uint8).copy()
        self.h = lib().ok_model_create(self._blob.ctypes.data, len(self._blob))
        if not self.h:
            raise RuntimeError(lib().ok_last_error().decode())
        self.S = lib().ok_snapshot_size(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().ok_model_destroy(self.h)
            self.h = None


class Data:
    def __init__(self, model: Model):
        self.model = model
        self.h = lib().ok_data_create(model.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().ok_data_destroy(self.h)
            self.h = None

    def load(self, snap):
        snap = np.ascontiguousarray(snap, dtype=np.float64)
        assert snap.size == self.model.S
        lib().ok_data_load_snapshot(self.h, _d(snap))

    def save(self):
        out = np.zeros(self.model.S)
        lib().ok_data_save_snapshot(self.h, _d(out))
        return out

    def step(self, skipstage=0, iterations=-1, tolerance=0.0):
        lib().ok_data_step(self.h, skipstage, iterations, tolerance)

    def forward(self, skipstage=0, iterations=-1, tolerance=0.0):
        lib().ok_data_forward(self.h, skipstage, iterations, tolerance)

    def get(self, name, cap=65536):
        out = np.zeros(cap)
        c = lib().ok_data_get(self.h, name.encode(), _d(out), cap)
        if c < 0:
            raise KeyError(name)
        return out[:c].copy()


def dynamics_derivs(model, task, snap, cols, central=True, eps=1e-6, iterations=5, tolerance=0.0):
    nx = 2 * task.n
    A = np.zeros((nx, nx), order="F")
    B = np.zeros((nx, max(task.m, 1)), order="F")
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    snap = np.ascontiguousarray(snap, dtype=np.float64)
    d = task.desc()
    rc = lib().ok_dynamics_derivs(model.h, C.byref(d), _d(snap), _i(cols), len(cols), int(central), eps, iterations,
                                  tolerance, _d(A), _d(B))
    if rc:
        raise RuntimeError(lib().ok_last_error().decode())
    return A, B[:, :task.m]


def flops_dynamics_derivs(model, task, snap, cols, iterations=5, tolerance=0.0):
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    snap = np.ascontiguousarray(snap, dtype=np.float64)
    d = task.desc()
    return int(lib().ok_flops_dynamics_derivs(model.h, C.byref(d), _d(snap), _i(cols), len(cols), iterations, tolerance))


def residual_derivs(model, task, snap, central=True, eps=1e-6):
    rx = np.zeros((task.nr, 2 * task.n), order="F")
    ru = np.zeros((task.nr, max(task.m, 1)), order="F")
    snap = np.ascontiguousarray(snap, dtype=np.float64)
    d = task.desc()
    rc = lib().ok_residual_derivs(model.h, C.byref(d), _d(snap), int(central), eps, _d(rx), _d(ru))
    if rc:
        raise RuntimeError(lib().ok_last_error().decode())
    return rx, ru[:, :task.m]


def init_sequence_ee(model, task, start, goal_xy, setup_horizon, T):
    """GenTestingData.cpp:177-220 for one BoxSweep (setup_horizon 0) / TwoDPushing (1000) trial:
    returns (snapshot0, U[T][m], U_setup[setup_horizon][m], warning bits)."""
    start = np.ascontiguousarray(start, dtype=np.float64)
    goal = np.ascontiguousarray(goal_xy, dtype=np.float64)
    snap0 = np.zeros(model.S)
    U = np.zeros((T, task.m))
    Us = np.zeros((max(setup_horizon, 1), task.m))
    d = task.desc()
    rc = lib().ok_init_sequence_ee(model.h, C.byref(d), _d(start), _d(goal), setup_horizon, T, _d(snap0), _d(U), _d(Us))
    if rc < 0:
        raise RuntimeError(lib().ok_last_error().decode())
    return snap0, U, Us[:setup_horizon], rc


def keypoints(method, X, n, T, dt, min_N, max_N, jerk_thr, vel_thr):
    X = np.ascontiguousarray(X, dtype=np.float64)
    jerk_thr = np.ascontiguousarray(jerk_thr, dtype=np.float64)
    vel_thr = np.ascontiguousarray(vel_thr, dtype=np.float64)
    counts = np.zeros(T, dtype=np.int32)
    idx = np.zeros(T * 2 * n + 8, dtype=np.int32)
    tot = lib().ok_keypoints(method, _d(X), n, T, dt, min_N, max_N, _d(jerk_thr), _d(vel_thr), _i(counts), _i(idx), len(idx))
    if tot < 0:
        raise RuntimeError("ok_keypoints failed")
    rows, p = [], 0
    for t in range(T):
        rows.append(idx[p:p + counts[t]].tolist())
        p += counts[t]
    return rows


def pack_rows(rows):
    counts = np.array([len(r) for r in rows], dtype=np.int32)
    idx = np.array([v for r in rows for v in r] + [0], dtype=np.int32)
    return counts, idx


def interpolate(rows, n, m, A, B):
    """A: (T, nx, nx) / B: (T, nx, m) arrays of column-major blocks, modified in place."""
    counts, idx = pack_rows(rows)
    lib().ok_interpolate(_i(counts), _i(idx), len(rows), n, m, _d(A), _d(B))


def percentages(rows, n):
    counts, idx = pack_rows(rows)
    pct = np.zeros(n)
    lib().ok_percentages(_i(counts), _i(idx), len(rows), n, _d(pct))
    return pct


class ILQR:
    def __init__(self, model: Model, task: Task, T, threads=0):
        self.model, self.task, self.T = model, task, T
        self._desc = task.desc()
        self.h = lib().ok_ilqr_create(model.h, C.byref(self._desc), T, threads)
        if not self.h:
            raise RuntimeError(lib().ok_last_error().decode())

    def __del__(self):
        if getattr(self, "h", None):
            lib().ok_ilqr_destroy(self.h)
            self.h = None

    def set_lambda(self, lam):
        lib().ok_ilqr_set_lambda(self.h, lam)

    def set_filter(self, method, low_pass_a=0.25, fir=(0.1, 0.15, 0.5, 0.15, 0.1)):
        """Optimiser::filteringMethod: 0 none, 1 low_pass, 2 FIR."""
        fir = np.ascontiguousarray(fir, dtype=np.float64)
        lib().ok_ilqr_set_filter(self.h, int(method), float(low_pass_a), _d(fir), len(fir))

    def filter(self):
        lib().ok_ilqr_filter(self.h)

    def set_A(self, A):
        A = np.ascontiguousarray(A, dtype=np.float64)
        lib().ok_ilqr_set_A(self.h, _d(A))

    def rollout(self, snap0, U):
        snap0 = np.ascontiguousarray(snap0, dtype=np.float64)
        U = np.ascontiguousarray(U, dtype=np.float64)
        return lib().ok_ilqr_rollout(self.h, _d(snap0), _d(U))

    def generate_keypoints(self):
        lib().ok_ilqr_generate_keypoints(self.h)

    def dynamics_derivs(self):
        lib().ok_ilqr_dynamics_derivs(self.h)

    def generate_derivatives(self):
        lib().ok_ilqr_generate_derivatives(self.h)

    def backward(self):
        return bool(lib().ok_ilqr_backward(self.h))

    def forward(self, thread_id, alpha):
        return lib().ok_ilqr_forward(self.h, thread_id, alpha)

    def save_best(self, thread_id, cost):
        lib().ok_ilqr_save_best(self.h, thread_id, cost)

    def iteration(self):
        le = C.c_int(0)
        conv = lib().ok_ilqr_iteration(self.h, C.byref(le))
        return bool(conv), bool(le.value)

    def optimise(self, snap0, U, max_it, min_it):
        snap0 = np.ascontiguousarray(snap0, dtype=np.float64)
        U = np.ascontiguousarray(U, dtype=np.float64)
        out = np.zeros((self.T, self.task.m))
        lib().ok_ilqr_optimise(self.h, _d(snap0), _d(U), max_it, min_it, _d(out))
        return out

    def get(self, name):
        cap = 1 << 16
        while True:
            out = np.zeros(cap)
            c = lib().ok_ilqr_get(self.h, name.encode(), _d(out), cap)
            if c <= -1000000:
                raise KeyError(name)
            if c < 0:
                cap = -c
                continue
            return out[:c].copy()

    def keypoints(self):
        n = self.task.n
        counts = np.zeros(self.T, dtype=np.int32)
        idx = np.zeros(self.T * 2 * n + 8, dtype=np.int32)
        tot = lib().ok_ilqr_get_keypoints(self.h, _i(counts), _i(idx), len(idx))
        assert tot >= 0
        rows, p = [], 0
        for t in range(self.T):
            rows.append(idx[p:p + counts[t]].tolist())
            p += counts[t]
        return rows

    def scalars(self):
        v = self.get("scalars")
        return dict(zip(["lambda", "delta_J", "old_cost", "new_cost", "initial_cost", "cost_reduction",
                         "num_iterations", "fd_sims"], v))
