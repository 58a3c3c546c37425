"""CPU: independent checks of the model compiler (csrc/mjcf_compile.cpp).  The compiled blob feeds BOTH the device and the
oracle, so device-vs-oracle tests cannot see a wrong mesh inertia, default-class resolution or invweight0 (VERDICT r1,
weak 1-iii).  Here the numbers are recomputed in numpy from the raw asset files with different formulas:
  * Panda link inertial properties from the collision STL: signed tetrahedra about the ORIGIN (divergence theorem) instead
    of the compiler's absolute tetrahedra about the vertex centroid; mass from the MJCF attribute;
  * dof_invweight0 from a dense numpy inverse of the mass matrix the ORACLE's composite-rigid-body pass builds at qpos0
    (the compiler builds its own by the Jacobian method and inverts by Cholesky);
  * body_invweight0 of free bodies in closed form (1 / m and mean 1 / I);
  * default-class resolution on the attributes the dynamics use (damping, armature, condim, contype / conaffinity)."""
import os
import struct
import xml.etree.ElementTree as ET

import numpy as np
import pytest

from helpers import make, oracle_model
from oracle import oracle
from trajoptkp_b200 import capi, tasks

MESHDIR = os.path.join(tasks.ASSETS, "models", "franka")


def _read_stl(path):
    raw = open(path, "rb").read()
    n = struct.unpack("<I", raw[80:84])[0]
    assert len(raw) == 84 + 50 * n, "binary STL expected"
    tri = np.zeros((n, 3, 3))
    for i in range(n):
        v = struct.unpack("<12f", raw[84 + 50 * i:84 + 50 * i + 48])
        tri[i] = np.array(v[3:]).reshape(3, 3)
    return tri


def _mass_properties(tri):
    """Volume, centre of mass and inertia tensor about it at unit density: signed tetrahedra (origin, a, b, c)."""
    a, b, c = tri[:, 0], tri[:, 1], tri[:, 2]
    vol6 = np.einsum("ij,ij->i", a, np.cross(b, c))
    if vol6.sum() < 0:
        vol6 = -vol6
    vol = vol6.sum() / 6.0
    com = ((a + b + c) / 4.0 * (vol6 / 6.0)[:, None]).sum(0) / vol
    # covariance of a tetrahedron with one vertex at the origin: V / 20 * (sum_i p_i p_i^T + s s^T), s = a + b + c
    s = a + b + c
    C = np.zeros((3, 3))
    for p in (a, b, c, s):
        C += np.einsum("i,ij,ik->jk", vol6 / 6.0 / 20.0, p, p)
    C -= vol * np.outer(com, com)
    return vol, com, np.trace(C) * np.eye(3) - C


def test_panda_link_inertias_from_the_stl_files():
    tc = make("reaching")
    m = tc.model
    root = ET.parse(tc.model_path).getroot()
    meshfile = {e.get("name"): e.get("file") for e in root.iter("mesh")}
    mass_all, inertia_all, ipos_all = m.real("body_mass"), m.real("body_inertia").reshape(-1, 3), m.real("body_ipos").reshape(-1, 3)
    checked = 0
    for body in root.iter("body"):
        cols = [g for g in body.findall("geom") if g.get("class") == "panda_col" and g.get("mass") and g.get("mesh") in meshfile]
        if len(cols) != 1 or body.find("inertial") is not None:
            continue
        others = [g for g in body.findall("geom") if g is not cols[0] and g.get("class") != "panda_viz"]
        if others:
            continue
        g = cols[0]
        if g.get("quat") is not None or g.get("euler") is not None:
            continue
        gpos = np.array([float(x) for x in g.get("pos").split()]) if g.get("pos") else np.zeros(3)
        mass = float(g.get("mass"))
        vol, com, I = _mass_properties(_read_stl(os.path.join(MESHDIR, meshfile[g.get("mesh")])))
        ev = np.sort(np.linalg.eigvalsh(I)) * mass / vol
        bid = m.name2id(capi.OBJ_BODY, body.get("name"))
        assert abs(mass_all[bid] - mass) < 1e-12, body.get("name")            # visual geoms carry no mass
        assert np.allclose(np.sort(inertia_all[bid]), ev, rtol=1e-8, atol=0), (body.get("name"), inertia_all[bid], ev)
        assert np.allclose(ipos_all[bid], gpos + com, rtol=0, atol=1e-10), (body.get("name"), ipos_all[bid], gpos + com)
        checked += 1
    assert checked >= 8  # link0 .. link7


@pytest.mark.parametrize("name", ["walker_run", "reaching", "box_sweep", "push_lcl"])
def test_dof_invweight0_is_the_diagonal_of_the_inverse_mass_matrix(built, name):
    tc = make(name)
    m = tc.model
    d = oracle.Data(oracle_model(tc))
    d.load(m.snapshot(qpos=m.real("qpos0")))
    d.forward()
    M = d.get("qM").reshape(m.nv, m.nv)
    assert np.allclose(M, M.T)
    Minv = np.linalg.inv(M)
    want = np.diag(Minv).copy()
    jt, da = m.int("jnt_type"), m.int("jnt_dofadr")
    for j in range(m.njnt):
        if jt[j] == 0:  # free joint: MuJoCo averages the translational and the rotational triple
            want[da[j]:da[j] + 3] = want[da[j]:da[j] + 3].mean()
            want[da[j] + 3:da[j] + 6] = want[da[j] + 3:da[j] + 6].mean()
    got = m.real("dof_invweight0")
    assert np.allclose(got, want, rtol=1e-9, atol=0), np.abs(got / want - 1).max()
    # stat.meaninertia = mean diagonal of M at qpos0 (the solver's cost scale)
    assert abs(m.real("opt_real")[7] - np.diag(M).mean()) <= 1e-10 * np.diag(M).mean()


def test_body_invweight0_of_free_bodies_in_closed_form():
    tc = make("push_lcl")
    m = tc.model
    mass, inertia = m.real("body_mass"), m.real("body_inertia").reshape(-1, 3)
    iw = m.real("body_invweight0").reshape(-1, 2)
    jt, jb = m.int("jnt_type"), m.int("jnt_bodyid")
    n = 0
    for j in range(m.njnt):
        if jt[j] != 0:
            continue
        b = int(jb[j])
        assert abs(iw[b, 0] - 1.0 / mass[b]) <= 1e-9 / mass[b], (b, iw[b], mass[b])
        assert abs(iw[b, 1] - np.mean(1.0 / inertia[b])) <= 1e-9 * np.mean(1.0 / inertia[b])
        n += 1
    assert n >= 5


def test_default_classes_reach_the_dynamics_attributes():
    tc = make("box_sweep")
    m = tc.model
    damping, armature = m.real("dof_damping"), m.real("dof_armature")
    da = m.int("jnt_dofadr")
    for k in range(1, 5):   # panda_arm: damping 100
        assert damping[da[m.name2id(capi.OBJ_JOINT, "panda0_joint%d" % k)]] == 100.0
    for k in range(5, 8):   # panda_forearm: damping 10
        assert damping[da[m.name2id(capi.OBJ_JOINT, "panda0_joint%d" % k)]] == 10.0
    f = da[m.name2id(capi.OBJ_JOINT, "panda0_finger_joint1")]
    assert damping[f] == 100.0 and armature[f] == 5.0   # panda_finger
    condim, gtype = m.int("geom_condim"), m.int("geom_type")
    assert set(condim.tolist()) <= {1, 3}               # the load-time check admits nothing else
    rng = m.real("jnt_range").reshape(-1, 2)
    j4 = m.name2id(capi.OBJ_JOINT, "panda0_joint4")
    assert np.allclose(rng[j4], [-3.0718, -0.4]) and m.int("jnt_limited")[j4] == 1   # radians (compiler angle="radian")
