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
TASK_ACROBOT, TASK_WALKER = 0, 1
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
                ("iterative_error_threshold", C.c_double)]


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
    L.tokp_step_worlds.argtypes = [vp, _dp, _dp, C.c_int, C.c_int, C.c_int, C.c_double, _ip]
    L.tokp_launch_count.restype = C.c_int64
    if hasattr(L, "tokp_traj_create"):
        L.tokp_traj_create.argtypes = [vp, C.POINTER(TaskDesc), C.c_int, C.c_int, C.POINTER(vp)]
        L.tokp_traj_destroy.argtypes = [vp]
        L.tokp_traj_set_targets.argtypes = [vp, C.c_int, _dp]
        L.tokp_traj_set_keypoint_method.argtypes = [vp, C.c_int, C.c_int, C.c_int, _dp, _dp, C.c_double]
        L.tokp_traj_set_snapshot.argtypes = [vp, C.c_int, C.c_int, _dp]
        L.tokp_traj_get_snapshot.argtypes = [vp, C.c_int, C.c_int, _dp]
        L.tokp_rollout.argtypes = [vp, _dp, _dp, _dp]
        L.tokp_keypoints.argtypes = [vp, _ip, _ip, _dp]
        L.tokp_traj_set_keypoints.argtypes = [vp, C.c_int, _ip, _ip]
        L.tokp_dynamics_derivs.argtypes = [vp, C.c_int, C.c_double]
        L.tokp_dynamics_derivs_cols.argtypes = [vp, C.c_int, C.c_int, _ip, C.c_int, C.c_int, C.c_double, _dp, C.c_int,
                                                _dp, C.c_int]
        L.tokp_interpolate.argtypes = [vp]
        L.tokp_cost_derivs.argtypes = [vp, C.c_int, C.c_double]
        L.tokp_generate_derivatives.argtypes = [vp]
        L.tokp_backward_pass.argtypes = [vp, _dp, _ip, _dp]
        L.tokp_forward_pass.argtypes = [vp, _dp, C.c_int, _dp]
        L.tokp_accept.argtypes = [vp, C.c_int, C.c_int]
        L.tokp_optimise.argtypes = [vp, _dp, _dp, C.c_int, C.c_int, _dp, _dp, _dp]
        L.tokp_traj_get.argtypes = [vp, C.c_int, C.c_char_p, _dp, C.c_int]
        L.tokp_traj_set.argtypes = [vp, C.c_int, C.c_char_p, _dp, C.c_int]
        L.tokp_traj_counters.argtypes = [vp, C.POINTER(C.c_int64)]
        L.tokp_timers.argtypes = [vp, _dp]
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
