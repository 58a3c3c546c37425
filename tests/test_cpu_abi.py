"""CPU tests of the drop-in boundary: libtokp.so loads, exports every symbol include/tokp.h declares, the host-side
model compiler produces the expected mjModel-equivalent fields for the shipped models, and every compute entry point
fails loudly (TOKP_ERR_CUDA) instead of falling back when there is no CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from trajoptkp_b200 import capi, tasks

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "tokp.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tokp_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built):
    L = capi.lib()
    syms = _declared_symbols()
    assert len(syms) >= 30
    missing = [s for s in syms if not hasattr(L, s)]
    assert not missing, missing
    assert L.tokp_version().decode()


def test_signatures_are_plain_c(built):
    """The boundary is a C ABI: no C++/torch types in the header."""
    src = open(os.path.join(ROOT, "include", "tokp.h")).read()
    assert 'extern "C"' in src
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    for bad in ("std::", "torch", "at::", "Eigen", "template"):
        assert bad not in src


@pytest.mark.parametrize("name,dims", [
    # nq, nv, nu, nbody, njnt, ngeom  
    ("acrobot", (2, 2, 1, 3, 2)),
    ("walker_run", (9, 9, 6, 8, 9)),
])
def test_model_compiler_sizes(built, name, dims):
    tc = tasks.TaskConfig(name)
    m = tc.model
    assert (m.nq, m.nv, m.nu, m.nbody, m.njnt) == dims
    assert m.S == 1 + m.nq + 4 * m.nv + 6 * m.nbody + m.nu  # the 8-field snapshot, MuJoCoHelper.cpp:732-741
    assert abs(m.real("opt_real")[0] - tc.timestep) < 1e-15   # YAML time step overrides the XML (MuJoCoHelper.cpp:895)


def test_walker_model_fields(built):
    m = tasks.TaskConfig("walker_run").model
    # two slide + seven hinge joints; the six leg joints are limited, ranges converted from degrees
    jt = m.int("jnt_type")
    assert list(jt[:3]) == [2, 2, 3] and all(jt[3:] == 3)
    lim = m.int("jnt_limited")
    assert list(lim) == [0, 0, 0, 1, 1, 1, 1, 1, 1]
    rng = m.real("jnt_range").reshape(-1, 2)
    assert np.allclose(rng[3], np.deg2rad([-20, 100]))
    assert np.allclose(rng[4], np.deg2rad([-150, 0]))
    assert np.allclose(rng[5], np.deg2rad([-45, 45]))
    # motors: gear 100/50/20 per leg, ctrlrange +-1
    assert np.allclose(m.real("actuator_gear").reshape(m.nu, -1)[:, 0], [100, 50, 20, 100, 50, 20])
    assert np.allclose(m.real("actuator_ctrlrange").reshape(-1, 2), [[-1, 1]] * 6)
    # kinematic tree: torso -> (thigh -> leg -> foot) x 2
    assert list(m.int("body_parentid")) == [0, 0, 1, 2, 3, 1, 5, 6]
    assert list(m.int("dof_parentid")) == [-1, 0, 1, 2, 3, 4, 2, 6, 7]
    # capsule masses at density 1000: m = rho * (pi r^2 L + 4/3 pi r^3)
    size = m.real("geom_size").reshape(-1, 3)
    gt = m.int("geom_type")
    gb = m.int("geom_bodyid")
    mass = m.real("body_mass")
    for b in range(1, m.nbody):
        tot = 0.0
        for gidx in range(m.ngeom):
            if gb[gidx] == b and gt[gidx] == 3:
                r, hl = size[gidx, 0], size[gidx, 1]
                tot += 1000.0 * (np.pi * r * r * 2 * hl + 4.0 / 3.0 * np.pi * r ** 3)
        assert abs(tot - mass[b]) < 1e-9 * tot, (b, tot, mass[b])
    assert np.allclose(m.real("dof_damping"), [0, 0, 0] + [0.1] * 6)
    assert np.allclose(m.real("dof_armature"), [0, 0, 0] + [0.01] * 6)


def test_acrobot_model_fields(built):
    m = tasks.TaskConfig("acrobot").model
    assert list(m.int("jnt_type")) == [3, 3]
    assert np.allclose(m.real("dof_damping"), [0.05, 0.05])
    assert np.allclose(m.real("actuator_gear").reshape(m.nu, -1)[:, 0], [100])
    assert list(m.int("body_parentid")) == [0, 0, 1]
    assert m.name2id(capi.OBJ_JOINT, "shoulder") == 0 and m.name2id(capi.OBJ_JOINT, "elbow") == 1
    assert m.name2id(capi.OBJ_JOINT, "no_such_joint") == -1


def test_model_blob_roundtrip_through_oracle_reader(built):
    from oracle import oracle
    tc = tasks.TaskConfig("walker_run")
    blob = tc.model.blob()
    assert blob[:8] == b"TOKPMDL1"
    om = oracle.Model(blob)
    assert om.S == tc.model.S


def test_task_index_maps(built):
    """ModelTranslator::InitModelTranslator ordering: robot joints in YAML order (ModelTranslator_Test.cpp:33-43)."""
    tc = tasks.TaskConfig("walker_run")
    assert (tc.n, tc.m, tc.nr) == (9, 6, 9)
    assert list(tc.dof_adr) == list(range(9)) and list(tc.qpos_adr) == list(range(9))
    assert list(tc.ctrl_adr) == list(range(6))
    ac = tasks.TaskConfig("acrobot")
    assert (ac.n, ac.m, ac.nr) == (2, 1, 5)
    start, targets = ac.load_trial(0)
    assert start.shape == (2,) and targets.shape == (5,)


def test_bad_model_reports_error(built):
    h = C.c_void_p()
    rc = capi.lib().tokp_model_compile(b"/nonexistent/model.xml", 0.0, C.byref(h))
    assert rc == capi_err("TOKP_ERR_MODEL")
    assert capi.lib().tokp_last_error()
    rc = capi.lib().tokp_model_compile_string(b"<mujoco><worldbody><body><joint type='weird'/></body></worldbody></mujoco>",
                                              b".", 0.0, C.byref(h))
    assert rc != 0


def capi_err(name):
    src = open(os.path.join(ROOT, "include", "tokp.h")).read()
    return int(re.search(name + r"\s*=\s*(\d+)", src).group(1))


def test_no_cpu_fallback_without_a_device(built):
    """Without a CUDA device the compute entry points return TOKP_ERR_CUDA; nothing is computed on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    tc = tasks.TaskConfig("acrobot")
    m = tc.model
    snaps = np.tile(tc.start_snapshot(), (2, 1))
    out = np.full_like(snaps, -7.0)
    st = np.zeros(2, dtype=np.int32)
    rc = capi.lib().tokp_step_worlds(m.h, capi._d(snaps), capi._d(out), 2, 1, -1, 0.0, capi._i(st))
    assert rc == capi_err("TOKP_ERR_CUDA")
    assert (out == -7.0).all()
    with pytest.raises(capi.TokpError):
        capi.Traj(tc, 10, 1)
