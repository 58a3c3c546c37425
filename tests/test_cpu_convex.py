"""Oracle convex collider (oracle/ok_convex.h: Minkowski Portal Refinement + support functions) against analytic
answers on shapes whose penetration is known in closed form, and the Franka reaching scene stepped on the CPU.
The collider restates MuJoCo's mjc_Convex / libccd ccdMPRPenetration (tolerance 1e-6): parity with MuJoCo's bits is
unpinned (MuJoCo absent), so these pins are geometric."""
import os
import struct

import numpy as np
import pytest

from helpers import make, oracle_model
from oracle import oracle
from trajoptkp_b200 import capi


def _write_box_stl(path, hx, hy, hz):
    """Binary STL of an axis-aligned box (12 outward-facing triangles)."""
    v = np.array([[sx * hx, sy * hy, sz * hz] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)])
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    tris = []
    for a, b, c, d in quads:
        tris += [(a, b, c), (a, c, d)]
    with open(path, "wb") as f:
        f.write(b"\0" * 80)
        f.write(struct.pack("<I", len(tris)))
        for t in tris:
            f.write(struct.pack("<3f", 0, 0, 0))
            for i in t:
                f.write(struct.pack("<3f", *v[i]))
            f.write(b"\0\0")


XML = """
<mujoco>
  <asset><mesh name="cube" file="cube.stl"/></asset>
  <worldbody>
    <geom name="floor" type="plane" size="5 5 1" contype="2" conaffinity="2"/>
    <body name="a" pos="0 0 1"><freejoint/><geom type="mesh" mesh="cube" mass="1" contype="1" conaffinity="1"/></body>
    <body name="b" pos="{bx} {by} {bz}"><freejoint/><geom type="{btype}" size="{bsize}" mass="1" contype="1" conaffinity="1"/></body>
    <body name="c" pos="3 0 {cz}"><freejoint/><geom type="mesh" mesh="cube" mass="1" contype="2" conaffinity="2"/></body>
  </worldbody>
</mujoco>
"""


def _scene(tmp_path, **kw):
    _write_box_stl(str(tmp_path / "cube.stl"), 0.1, 0.2, 0.3)
    args = dict(bx=0.19, by=0.0, bz=1.0, btype="box", bsize="0.1 0.1 0.1", cz=2.0)
    args.update(kw)
    (tmp_path / "m.xml").write_text(XML.format(**args))
    m = capi.Model(str(tmp_path / "m.xml"), 0.002)
    om = oracle.Model(m.blob())
    d = oracle.Data(om)
    d.load(m.snapshot())
    d.forward()
    return m, d


def test_mesh_compile_of_a_box(tmp_path):
    m, d = _scene(tmp_path)
    # mass 1 box of half-sizes (0.1, 0.2, 0.3): principal moments m/3 (b^2 + c^2) ..., sorted as eig3 orders them
    want = sorted([(0.2 ** 2 + 0.3 ** 2) / 3, (0.1 ** 2 + 0.3 ** 2) / 3, (0.1 ** 2 + 0.2 ** 2) / 3])
    got = sorted(m.real("body_inertia").reshape(-1, 3)[1])
    np.testing.assert_allclose(got, want, rtol=1e-6)  # STL stores float32
    np.testing.assert_allclose(m.real("body_ipos").reshape(-1, 3)[1], 0, atol=1e-9)
    assert m.int("geom_meshnum")[1] == 8


def test_mesh_box_face_penetration(tmp_path):
    """Cube mesh (half-width 0.1 along x) against a box whose face overlaps it by 0.01."""
    m, d = _scene(tmp_path)
    assert d.get("ncon")[0] == 1 and d.get("warning")[0] == 0
    np.testing.assert_allclose(d.get("contact_dist"), [-0.01], atol=2e-6)
    frame = d.get("contact_frame").reshape(-1, 9)[0]
    np.testing.assert_allclose(frame[:3], [-1, 0, 0], atol=1e-5)  # pairs are ordered by geom type: box -> mesh
    pos = d.get("contact_pos").reshape(-1, 3)[0]
    assert abs(pos[0] - 0.095) < 2e-6  # midway between the two overlapping faces


def test_mesh_sphere_and_separation(tmp_path):
    m, d = _scene(tmp_path, btype="sphere", bsize="0.05", bx=0.13)
    assert d.get("ncon")[0] == 1
    np.testing.assert_allclose(d.get("contact_dist"), [-0.02], atol=2e-6)
    np.testing.assert_allclose(d.get("contact_frame").reshape(-1, 9)[0][:3], [-1, 0, 0], atol=1e-4)  # sphere -> mesh
    m, d = _scene(tmp_path, btype="sphere", bsize="0.05", bx=0.151)
    assert d.get("ncon")[0] == 0
    # diagonal separation: bounding spheres overlap, the shapes do not
    m, d = _scene(tmp_path, btype="box", bsize="0.1 0.1 0.1", bx=0.21, by=0.31)
    assert d.get("ncon")[0] == 0


def test_plane_mesh_deepest_vertex(tmp_path):
    m, d = _scene(tmp_path, cz=0.295)
    assert d.get("ncon")[0] == 2  # mesh-box pair of the default scene + plane-mesh
    dist = d.get("contact_dist")
    assert np.isclose(dist, -0.005, atol=1e-7).any()


def test_reaching_scene_steps_on_the_cpu():
    tc = make("reaching")
    om = oracle_model(tc)
    d = oracle.Data(om)
    d.load(tc.start_snapshot())
    d.forward()
    assert d.get("warning")[0] == 0
    # the finger pads touch at the closed-gripper start configuration: one mesh-mesh contact near the hand
    assert d.get("ncon")[0] >= 1
    q0 = d.get("qpos").copy()
    for _ in range(50):
        d.step()
    q = d.get("qpos")
    assert np.isfinite(q).all() and d.get("warning")[0] == 0
    # the gravity-compensated target sphere does not move, the unactuated arm sags
    np.testing.assert_allclose(q[9:16], q0[9:16], atol=1e-9)
    assert abs(q[1] - q0[1]) > 1e-3
