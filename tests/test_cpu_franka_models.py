"""Franka-scene models (reaching / box_sweep / push_*): MJCF subset compiler with mesh assets.

Pins: model sizes of SURVEY.md Appendix A (nq / nv / nu / nbody read off the reference's XMLs), the link masses
the XMLs state, and the structure MuJoCo's compiler gives mesh geoms (frame at the centre of mass, principal axes,
visual-only meshes never opened).  No compute calls: runs without a GPU.
"""
import os

import numpy as np
import pytest

from trajoptkp_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FRANKA = os.path.join(ROOT, "assets", "models", "franka")

# (file, timestep of the task YAML, nq, nv, nu, nbody) -- SURVEY.md Appendix A
SIZES = [
    ("reaching_scene", 0.008, 16, 15, 7, 13),
    ("boxSweep", 0.008, 23, 21, 7, 14),
    ("cylinder_pushing", 0.008, 23, 21, 7, 14),
    ("cylinder_pushing_lowClutter", 0.008, 44, 39, 7, 17),
    ("cylinder_pushing_heavyClutter", 0.01, 72, 63, 7, 21),
]


@pytest.mark.parametrize("name,dt,nq,nv,nu,nbody", SIZES)
def test_model_sizes(name, dt, nq, nv, nu, nbody):
    m = capi.Model(os.path.join(FRANKA, name + ".xml"), dt)
    d = m.int("dims")
    assert (d[0], d[1], d[2], d[3]) == (nq, nv, nu, nbody)
    assert m.real("opt_real")[0] == dt
    # every collision mesh was loaded, every visual one was not
    gt, adr, num = m.int("geom_type"), m.int("geom_meshadr"), m.int("geom_meshnum")
    ct, ca = m.int("geom_contype"), m.int("geom_conaffinity")
    for g in range(len(gt)):
        if gt[g] != 7:
            assert adr[g] == -1
            continue
        if ct[g] == 0 and ca[g] == 0:
            assert adr[g] == -1 and num[g] == 0
        else:
            assert adr[g] >= 0 and num[g] >= 4
    assert d[7] == len(m.real("mesh_vert")) // 3


def test_panda_links():
    m = capi.Model(os.path.join(FRANKA, "reaching_scene.xml"), 0.008)
    mass = m.real("body_mass")
    # geom masses of chain0.xml, explicit <inertial> of the fingers, sphere of density 1000 for the target
    want = [0, 2.91242, 2.7063, 2.73046, 2.04104, 2.08129, 3.00049, 1.3235, 0.2, 0.81909, 0.0927059, 0.0927059,
            1000 * 4 / 3 * np.pi * 0.1 ** 3]
    np.testing.assert_allclose(mass, want, rtol=1e-12)
    inertia = m.real("body_inertia").reshape(-1, 3)
    assert (inertia[1:] > 0).all()
    # a solid body satisfies the triangle inequality on its principal moments
    for i in inertia[1:]:
        assert i[0] <= i[1] + i[2] + 1e-12 and i[1] <= i[0] + i[2] + 1e-12 and i[2] <= i[0] + i[1] + 1e-12
    assert m.real("body_gravcomp")[-1] == 1.0
    np.testing.assert_array_equal(m.real("dof_damping")[:9], [100, 100, 100, 100, 10, 10, 10, 100, 100])
    np.testing.assert_array_equal(m.real("dof_armature")[7:9], [5, 5])
    # motors with the torque limits of the task YAML
    np.testing.assert_array_equal(m.real("actuator_ctrlrange").reshape(-1, 2)[:, 1], [87, 87, 87, 87, 12, 12, 12])


def test_mesh_frame_is_principal():
    """Vertices are stored in the frame MuJoCo gives a mesh: box bound = geom_size, bounding radius = its norm; the
    vertex cloud of a convex link is centred near its centre of mass."""
    m = capi.Model(os.path.join(FRANKA, "reaching_scene.xml"), 0.008)
    gt, adr, num = m.int("geom_type"), m.int("geom_meshadr"), m.int("geom_meshnum")
    vert = m.real("mesh_vert").reshape(-1, 3)
    size = m.real("geom_size").reshape(-1, 3)
    rb = m.real("geom_rbound")
    seen = 0
    for g in range(len(gt)):
        if gt[g] != 7 or adr[g] < 0:
            continue
        v = vert[adr[g]:adr[g] + num[g]]
        np.testing.assert_allclose(np.abs(v).max(axis=0), size[g], rtol=1e-12)
        assert abs(rb[g] - np.linalg.norm(size[g])) < 1e-12
        assert np.linalg.norm(v, axis=1).max() <= rb[g] + 1e-12
        seen += 1
    assert seen == 11  # link0..7, hand, two fingers


def test_collision_pairs_follow_the_filters():
    m = capi.Model(os.path.join(FRANKA, "boxSweep.xml"), 0.008)
    g1, g2 = m.int("pair_geom1"), m.int("pair_geom2")
    body = m.int("geom_bodyid")
    ct, ca = m.int("geom_contype"), m.int("geom_conaffinity")
    weld = m.int("body_weldid")
    parent = m.int("body_parentid")
    for a, b in zip(g1, g2):
        assert (ct[a] & ca[b]) or (ct[b] & ca[a])
        wa, wb = weld[body[a]], weld[body[b]]
        assert wa != wb
        if wa and wb:
            assert weld[parent[wa]] != wb and weld[parent[wb]] != wa
