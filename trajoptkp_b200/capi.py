"""ctypes binding of libtokp.so (the C ABI declared in include/tokp.h).

This is plumbing for the Python harness (tests, bench, sweep driver); the product is the shared
library.  Loading fails loudly when the library has not been built: there is no fallback.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TOKP_LIB", os.path.join(_HERE, "_build", "libtokp.so"))

OBJ_BODY, OBJ_JOINT, OBJ_GEOM, OBJ_SITE, OBJ_ACTUATOR = 1, 3, 5, 6, 19
TASK_ACROBOT, TASK_WALKER, TASK_REACHING, TASK_BOXSWEEP, TASK_PUSH = 0, 1, 2, 3, 4
KP_SET_INTERVAL, KP_ADAPTIVE_JERK, KP_ADAPTIVE_ACCEL, KP_ITERATIVE_ERROR, KP_VELOCITY_CHANGE = range(5)
KP_NAMES = {"set_interval": 0, "setInterval": 0, "adaptive_jerk": 1, "adaptive_accel": 2, "iterative_error": 3,
            "velocity_change": 4}

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


class TaskDesc(C.Structure):
    _fields_ = [("kind", C.c_int), ("n", C.c_int), ("m", C.c_int),
                ("qpos_adr", _ip), ("dof_adr", _ip), ("ctrl_adr", _ip),
                ("num_resid", C.c_int),
                ("resid_target", _dp), ("resid_weight", _dp), ("resid_weight_terminal", _dp),
                ("kp_method", C.c_int), ("min_N", C.c_int), ("max_N", C.c_int),
                ("jerk_thresholds", _dp), ("velchange_thresholds", _dp),
                ("iterative_error_threshold", C.c_double),
                ("num_target", C.c_int), ("num_aux", C.c_int), ("aux", _ip)]


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def _i(a):
    return a.ctypes.data_as(_ip) if a is not None else None


_lib = None


class TokpError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise TokpError("libtokp.so not built (%s): run `python -c 'import __graft_entry__ as g; g.build()'`" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.tokp_last_error.restype = C.c_char_p
    L.tokp_version.restype = C.c_char_p
    L.tokp_model_compile.argtypes = [C.c_char_p, C.c_double, C.POINTER(vp)]
    L.tokp_model_compile_string.argtypes = [C.c_char_p, C.c_char_p, C.c_double, C.POINTER(vp)]
    L.tokp_model_destroy.argtypes = [vp]
    L.tokp_model_dims.argtypes = [vp, _ip]
    L.tokp_model_name2id.argtypes = [vp, C.c_int, C.c_char_p]
    L.tokp_model_blob.restype = C.c_int64
    L.tokp_model_blob.argtypes = [vp, vp, C.c_int64]
    L.tokp_model_get_real.argtypes = [vp, C.c_char_p, _dp, C.c_int]
    L.tokp_model_get_int.argtypes = [vp, C.c_char_p, C.POINTER(C.c_int32), C.c_int]
    L.tokp_snapshot_size.argtypes = [vp]
    L.tokp_model_body_pose.argtypes = [vp, _dp, C.c_int, _dp, _dp]
    L.tokp_step_worlds.argtypes = [vp, _dp, _dp, C.c_int, C.c_int, C.c_int, C.c_double, _ip]
    L.tokp_forward_worlds.argtypes = [vp, _dp, _dp, C.c_int, C.c_int, C.c_int, C.c_double, _ip]
    L.tokp_step_worlds_static.argtypes = [vp, _dp, _dp, C.c_int, C.c_int, C.c_int, C.c_double, _ip]
    L.tokp_launch_count.restype = C.c_int64
    if hasattr(L, "tokp_traj_create"):
        L.tokp_traj_create.argtypes = [vp, C.POINTER(TaskDesc), C.c_int, C.c_int, C.POINTER(vp)]
        L.tokp_traj_destroy.argtypes = [vp]
        L.tokp_traj_set_targets.argtypes = [vp, C.c_int, _dp]
        L.tokp_traj_set_keypoint_method.argtypes = [vp, C.c_int, C.c_int, C.c_int, _dp, _dp, C.c_double]
        L.tokp_traj_set_snapshot.argtypes = [vp, C.c_int, C.c_int, _dp]
        L.tokp_traj_get_snapshot.argtypes = [vp, C.c_int, C.c_int, _dp]
        L.tokp_rollout.argtypes = [vp, _dp, _dp, _dp]
        L.tokp_init_controls_gravcomp.argtypes = [vp, _dp, _dp, _ip]
        L.tokp_init_controls_ee.argtypes = [vp, _dp, _dp, C.c_int, C.c_int, _dp, _dp, _dp, _ip]
        L.tokp_keypoints.argtypes = [vp, _ip, _ip, _dp]
        L.tokp_traj_set_keypoints.argtypes = [vp, C.c_int, _ip, _ip]
        L.tokp_dynamics_derivs.argtypes = [vp, C.c_int, C.c_double]
        L.tokp_dynamics_derivs_cols.argtypes = [vp, C.c_int, C.c_int, _ip, C.c_int, C.c_int, C.c_double, _dp, C.c_int,
                                                _dp, C.c_int]
        L.tokp_interpolate.argtypes = [vp]
        L.tokp_cost_derivs.argtypes = [vp, C.c_int, C.c_double]
        L.tokp_generate_derivatives.argtypes = [vp]
        L.tokp_residual_derivs.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_double, _dp, _dp]
        L.tokp_traj_set_filter.argtypes = [vp, C.c_int, C.c_double, _dp, C.c_int]
        L.tokp_filter_dynamics.argtypes = [vp]
        L.tokp_backward_pass.argtypes = [vp, _dp, _ip, _dp]
        L.tokp_forward_pass.argtypes = [vp, _dp, C.c_int, _dp]
        L.tokp_accept.argtypes = [vp, C.c_int, C.c_int]
        L.tokp_optimise.argtypes = [vp, _dp, _dp, C.c_int, C.c_int, _dp, _dp, _dp]
        L.tokp_traj_get.argtypes = [vp, C.c_int, C.c_char_p, _dp, C.c_int]
        L.tokp_traj_set.argtypes = [vp, C.c_int, C.c_char_p, _dp, C.c_int]
        L.tokp_traj_status.argtypes = [vp, _ip, C.c_int]
        L.tokp_traj_set_option.argtypes = [vp, C.c_char_p, C.c_int]
        L.tokp_traj_counters.argtypes = [vp, C.POINTER(C.c_int64)]
        L.tokp_timers.argtypes = [vp, _dp]
        L.tokp_fd_kernel_ms.argtypes = [vp, _dp]
        L.tokp_measure_peaks.argtypes = [_dp]
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise TokpError("tokp error %d: %s" % (rc, lib().tokp_last_error().decode()))


class Model:
    """Compiled model handle (tokp_model)."""

    def __init__(self, mjcf_path, timestep=0.0):
        h = C.c_void_p()
        check(lib().tokp_model_compile(os.fsencode(mjcf_path), float(timestep), C.byref(h)))
        self.h = h
        dims = (C.c_int * 11)()
        check(lib().tokp_model_dims(self.h, dims))
        (self.nq, self.nv, self.nu, self.nbody, self.njnt, self.ngeom, self.npair, self.nmeshvert, self.maxcontact,
         self.maxefc, self.treedepth) = list(dims)
        self.S = lib().tokp_snapshot_size(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().tokp_model_destroy(self.h)
            self.h = None

    def name2id(self, objtype, name):
        return lib().tokp_model_name2id(self.h, objtype, name.encode())

    def blob(self):
        n = lib().tokp_model_blob(self.h, None, 0)
        buf = np.zeros(n, dtype=np.uint8)
        assert lib().tokp_model_blob(self.h, buf.ctypes.data, n) == n
        return buf.tobytes()

    def real(self, field):
        c = lib().tokp_model_get_real(self.h, field.encode(), None, 0)
        if c <= -1000000:
            raise KeyError(field)
        out = np.zeros(-c if c < 0 else c)
        if out.size:
            lib().tokp_model_get_real(self.h, field.encode(), _d(out), out.size)
        return out

    def int(self, field):
        c = lib().tokp_model_get_int(self.h, field.encode(), None, 0)
        if c <= -1000000:
            raise KeyError(field)
        out = np.zeros(-c if c < 0 else c, dtype=np.int32)
        if out.size:
            lib().tokp_model_get_int(self.h, field.encode(), out.ctypes.data_as(C.POINTER(C.c_int32)), out.size)
        return out

    # ---- snapshot helpers (layout of MuJoCoHelper::CpMjData) -------------------------------------
    def body_pose(self, qpos, body):
        """World pose (pos[3], quat[4]) of a body at qpos (host kinematics of the chain)."""
        q = np.ascontiguousarray(qpos, dtype=np.float64)
        assert q.size == self.nq
        pos, quat = np.zeros(3), np.zeros(4)
        check(lib().tokp_model_body_pose(self.h, _d(q), int(body), _d(pos), _d(quat)))
        return pos, quat

    def snapshot(self, qpos=None, qvel=None, ctrl=None, qacc=None, warm=None, time=0.0):
        s = np.zeros(self.S)
        s[0] = time
        s[1:1 + self.nq] = self.real("qpos0") if qpos is None else qpos
        o = 1 + self.nq
        if qvel is not None:
            s[o:o + self.nv] = qvel
        if qacc is not None:
            s[o + self.nv:o + 2 * self.nv] = qacc
        if warm is not None:
            s[o + 2 * self.nv:o + 3 * self.nv] = warm
        if ctrl is not None:
            s[self.S - self.nu:] = ctrl
        return s

    def split(self, s):
        o = 1 + self.nq
        nv = self.nv
        return dict(time=s[0], qpos=s[1:o], qvel=s[o:o + nv], qacc=s[o + nv:o + 2 * nv], warm=s[o + 2 * nv:o + 3 * nv],
                    ctrl=s[self.S - self.nu:])


class Traj:
    """Batch of B trajectories on the device (tokp_traj)."""

    def __init__(self, task_cfg, horizon, batch=1):
        self.tc, self.model, self.T, self.B = task_cfg, task_cfg.model, horizon, batch
        self.n, self.m, self.nr = task_cfg.n, task_cfg.m, task_cfg.nr
        self.nx = 2 * self.n
        self._desc = task_cfg.desc()
        h = C.c_void_p()
        check(lib().tokp_traj_create(self.model.h, C.byref(self._desc), horizon, batch, C.byref(h)))
        self.h = h

    def __del__(self):
        if getattr(self, "h", None):
            lib().tokp_traj_destroy(self.h)
            self.h = None

    def set_targets(self, b, targets):
        t = np.ascontiguousarray(targets, dtype=np.float64)
        assert t.size == (self._desc.num_target or self.nr)
        check(lib().tokp_traj_set_targets(self.h, b, _d(t)))

    def set_keypoint_method(self, method, min_N, max_N, jerk_thr=None, vel_thr=None, iter_thr=0.0):
        j = None if jerk_thr is None else np.ascontiguousarray(jerk_thr, dtype=np.float64)
        v = None if vel_thr is None else np.ascontiguousarray(vel_thr, dtype=np.float64)
        check(lib().tokp_traj_set_keypoint_method(self.h, method, min_N, max_N, _d(j), _d(v), float(iter_thr)))

    def set_snapshot(self, b, t, snap):
        s = np.ascontiguousarray(snap, dtype=np.float64)
        check(lib().tokp_traj_set_snapshot(self.h, b, t, _d(s)))

    def get_snapshot(self, b, t):
        s = np.zeros(self.model.S)
        check(lib().tokp_traj_get_snapshot(self.h, b, t, _d(s)))
        return s

    def rollout(self, snap0, U):
        s = np.ascontiguousarray(np.broadcast_to(np.asarray(snap0, dtype=np.float64), (self.B, self.model.S)))
        U = np.ascontiguousarray(np.broadcast_to(np.asarray(U, dtype=np.float64), (self.B, self.T, self.m)))
        cost = np.zeros(self.B)
        check(lib().tokp_rollout(self.h, _d(s), _d(U), _d(cost)))
        return cost

    def init_controls_gravcomp(self, snap0):
        """pandaReaching::CreateInitOptimisationControls on the device: U[B][T][m]."""
        s = np.ascontiguousarray(np.broadcast_to(np.asarray(snap0, dtype=np.float64), (self.B, self.model.S)))
        U = np.zeros((self.B, self.T, self.m))
        st = np.zeros(self.B, dtype=np.int32)
        check(lib().tokp_init_controls_gravcomp(self.h, _d(s), _d(U), _i(st)))
        return U, st

    def init_controls_ee(self, snap_start, goal_xy, pre_steps=1, setup_horizon=0, want_setup=False):
        """GenTestingData.cpp:177-220 for BoxSweep (setup_horizon 0) / TwoDPushing (1000) trials on the device:
        returns (snapshot0[B][S], U[B][T][m], status[B]) (+ U_setup[B][setup_horizon][m] when want_setup)."""
        s = np.ascontiguousarray(np.broadcast_to(np.asarray(snap_start, dtype=np.float64), (self.B, self.model.S)))
        g = np.ascontiguousarray(np.broadcast_to(np.asarray(goal_xy, dtype=np.float64), (self.B, 2)))
        snap0 = np.zeros((self.B, self.model.S))
        U = np.zeros((self.B, self.T, self.m))
        Us = np.zeros((self.B, setup_horizon, self.m)) if (want_setup and setup_horizon) else None
        st = np.zeros(self.B, dtype=np.int32)
        check(lib().tokp_init_controls_ee(self.h, _d(s), _d(g), pre_steps, setup_horizon, _d(snap0), _d(U), _d(Us), _i(st)))
        return (snap0, U, st, Us) if want_setup else (snap0, U, st)

    def set_option(self, name, value):
        check(lib().tokp_traj_set_option(self.h, name.encode(), int(value)))

    def status(self, clear=False):
        """[B][2] sticky step-warning bits of the FD worlds / rollout worlds of each trajectory (tokp_traj_status)."""
        out = np.zeros((self.B, 2), dtype=np.int32)
        check(lib().tokp_traj_status(self.h, _i(out), int(clear)))
        return out

    def keypoints(self):
        counts = np.zeros((self.B, self.T), dtype=np.int32)
        idx = np.zeros((self.B, self.T, 2 * self.n), dtype=np.int32)
        pct = np.zeros((self.B, self.n))
        check(lib().tokp_keypoints(self.h, _i(counts), _i(idx), _d(pct)))
        rows = [[idx[b, t, :counts[b, t]].tolist() for t in range(self.T)] for b in range(self.B)]
        return rows, pct

    def set_keypoints(self, b, rows):
        counts = np.array([len(r) for r in rows], dtype=np.int32)
        idx = np.array([v for r in rows for v in r] + [0], dtype=np.int32)
        check(lib().tokp_traj_set_keypoints(self.h, b, _i(counts), _i(idx)))

    def dynamics_derivs(self, central=True, eps=1e-6):
        check(lib().tokp_dynamics_derivs(self.h, int(central), eps))

    def dynamics_derivs_cols(self, b, t, cols, central=True, eps=1e-6, A=None, B=None):
        nx = self.nx
        A = np.zeros((nx, nx), order="F") if A is None else A
        Bm = np.zeros((nx, max(self.m, 1)), order="F") if B is None else B
        cols = np.ascontiguousarray(cols, dtype=np.int32)
        check(lib().tokp_dynamics_derivs_cols(self.h, b, t, _i(cols), len(cols), int(central), eps, _d(A), nx, _d(Bm), nx))
        return A, Bm[:, :self.m]

    def interpolate(self):
        check(lib().tokp_interpolate(self.h))

    def cost_derivs(self, central=True, eps=1e-6):
        check(lib().tokp_cost_derivs(self.h, int(central), eps))

    def generate_derivatives(self):
        check(lib().tokp_generate_derivatives(self.h))

    def residual_derivs(self, b, t, central=True, eps=1e-6):
        """Differentiator::ResidualDerivatives at (b, t): r_x [nx][nr], r_u [m][nr]."""
        rx = np.zeros((self.nx, self.nr))
        ru = np.zeros((max(self.m, 1), self.nr))
        check(lib().tokp_residual_derivs(self.h, b, t, int(central), eps, _d(rx), _d(ru)))
        return rx, ru[:self.m]

    FILTERS = {"none": 0, "low_pass": 1, "FIR": 2}

    def set_filter(self, method, low_pass_a=0.25, fir=None):
        """Optimiser::filteringMethod ("none" | "low_pass" | "FIR"), lowPassACoefficient, setFIRFilter."""
        code = self.FILTERS[method] if isinstance(method, str) else int(method)
        if fir is None:
            check(lib().tokp_traj_set_filter(self.h, code, float(low_pass_a), None, 0))
        else:
            f = np.ascontiguousarray(fir, dtype=np.float64)
            check(lib().tokp_traj_set_filter(self.h, code, float(low_pass_a), _d(f), len(f)))

    def filter_dynamics(self):
        check(lib().tokp_filter_dynamics(self.h))

    def backward_pass(self, lam):
        lam = np.ascontiguousarray(np.broadcast_to(np.asarray(lam, dtype=np.float64), (self.B,)))
        ok = np.zeros(self.B, dtype=np.int32)
        dJ = np.zeros(self.B)
        check(lib().tokp_backward_pass(self.h, _d(lam), _i(ok), _d(dJ)))
        return ok, dJ

    def forward_pass(self, alphas):
        a = np.ascontiguousarray(alphas, dtype=np.float64)
        costs = np.zeros((self.B, len(a)))
        check(lib().tokp_forward_pass(self.h, _d(a), len(a), _d(costs)))
        return costs

    def accept(self, b, alpha_idx):
        check(lib().tokp_accept(self.h, b, alpha_idx))

    def fd_kernel_ms(self):
        t = np.zeros(3)
        check(lib().tokp_fd_kernel_ms(self.h, _d(t)))
        return t

    def optimise_resident(self, max_it, min_it, lam=None, want_stats=True):
        """tokp_optimise on the device-resident inputs of the previous call (no bulk host<->device copies)."""
        lamv = None if lam is None else np.ascontiguousarray(np.broadcast_to(np.asarray(lam, dtype=np.float64), (self.B,)))
        stats = np.zeros((self.B, 8)) if want_stats else None
        check(lib().tokp_optimise(self.h, None, None, max_it, min_it, _d(lamv), None, _d(stats)))
        return stats

    def optimise(self, snap0, U, max_it, min_it, lam=None):
        s = np.ascontiguousarray(np.broadcast_to(np.asarray(snap0, dtype=np.float64), (self.B, self.model.S)))
        U = np.ascontiguousarray(np.broadcast_to(np.asarray(U, dtype=np.float64), (self.B, self.T, self.m)))
        lamv = None if lam is None else np.ascontiguousarray(np.broadcast_to(np.asarray(lam, dtype=np.float64), (self.B,)))
        Uo = np.zeros((self.B, self.T, self.m))
        stats = np.zeros((self.B, 8))
        check(lib().tokp_optimise(self.h, _d(s), _d(U), max_it, min_it, _d(lamv), _d(Uo), _d(stats)))
        return Uo, stats

    def get(self, b, name):
        c = lib().tokp_traj_get(self.h, b, name.encode(), None, 0)
        if c <= -1000000:
            raise KeyError(name)
        out = np.zeros(-c if c < 0 else c)
        got = lib().tokp_traj_get(self.h, b, name.encode(), _d(out), out.size)
        assert got == out.size, (name, got)
        return out

    def set(self, b, name, arr):
        a = np.ascontiguousarray(arr, dtype=np.float64).ravel()
        check(lib().tokp_traj_set(self.h, b, name.encode(), _d(a), a.size))

    def counters(self):
        c = (C.c_int64 * 4)()
        check(lib().tokp_traj_counters(self.h, c))
        return list(c)

    def timers(self):
        t = np.zeros(3)
        check(lib().tokp_timers(self.h, _d(t)))
        return t


def measure_peaks():
    out = np.zeros(2)
    check(lib().tokp_measure_peaks(_d(out)))
    return {"fp64_tflops": float(out[0]), "copy_gbs": float(out[1])}
