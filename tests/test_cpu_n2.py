"""CPU: the oracle's restatement of the pushing tasks' initial-control generators (oracle/ok_init.h;
PushBaseClass.cpp:8-248, BoxSweep.cpp:63-99, TwoDPushing.cpp:225-290, GenTestingData.cpp:177-220) against independent
numpy restatements of its pieces and against what the reference's results require of it (the gripper reaches the
object, iLQR then makes progress), plus the host pieces of the product that feed it (InitialiseSystemToStartState,
tokp_model_body_pose)."""
import numpy as np
import pytest

from helpers import make, oracle_model, oracle_task
from oracle import oracle
from trajoptkp_b200 import capi

REF_PI = 3.14152  # include/StdInclude.h:13


def _play(om, tc, snap0, U):
    d = oracle.Data(om)
    d.load(snap0)
    ncon, ee = 0, []
    gid = tc.model.name2id(capi.OBJ_BODY, "franka_gripper")
    for t in range(len(U)):
        s = d.save()
        s[tc.model.S - tc.model.nu + tc.ctrl_adr] = U[t]
        d.load(s)
        d.step()
        ncon = max(ncon, int(d.get("ncon")[0]))
        ee.append(d.get("xpos")[3 * gid:3 * gid + 3])
    return ncon, np.array(ee), d


def test_init_snapshot_is_initialise_system_to_start_state(built):
    """ModelTranslator.cpp:1711-1781: everything zero, goal visual placed, robot joints and every state-vector body at
    the filed start pose with an identity quaternion for zero angles."""
    tc = make("box_sweep")
    m = tc.model
    snap, goal = tc.init_snapshot(0)
    f = m.split(snap)
    joints, bodies, rest = tc.trial_row(0)
    assert np.array_equal(goal, rest[:2])
    assert np.array_equal(f["qpos"][tc.qpos_adr[:7]], joints)
    assert not f["qvel"].any() and not f["ctrl"].any() and f["time"] == 0.0
    qa = int(m.int("jnt_qposadr")[int(m.int("body_jntadr")[m.name2id(capi.OBJ_BODY, "goal")])])
    assert np.array_equal(f["qpos"][qa:qa + 7], [0.65, 0.0, 0.16, 1.0, 0.0, 0.0, 0.0])
    qd = int(m.int("jnt_qposadr")[int(m.int("body_jntadr")[m.name2id(capi.OBJ_BODY, "display_goal")])])
    assert np.array_equal(f["qpos"][qd:qd + 7], [goal[0], goal[1], 0.0, 1.0, 0.0, 0.0, 0.0])  # BoxSweep.cpp:52-61
    # fingers: zeroed, not qpos0
    fa = int(m.int("jnt_qposadr")[m.name2id(capi.OBJ_JOINT, "panda0_finger_joint1")])
    assert f["qpos"][fa] == 0.0


def test_body_pose_matches_oracle_kinematics(built):
    tc = make("reaching")
    m = tc.model
    om = oracle_model(tc)
    rng = np.random.default_rng(3)
    d = oracle.Data(om)
    for _ in range(5):
        snap, _ = tc.init_snapshot(int(rng.integers(0, 100)))
        f = m.split(snap)
        f["qpos"][tc.qpos_adr[:7]] += rng.uniform(-0.3, 0.3, 7)
        snap = m.snapshot(qpos=f["qpos"])
        d.load(snap)
        d.forward()
        for name in ("franka_gripper", "panda0_link4", "target", "panda0_leftfinger"):
            b = m.name2id(capi.OBJ_BODY, name)
            pos, quat = m.body_pose(f["qpos"], b)
            assert np.abs(pos - d.get("xpos")[3 * b:3 * b + 3]).max() < 1e-14
            assert np.abs(quat - d.get("xquat")[4 * b:4 * b + 4]).max() < 1e-14


def test_reaching_goal_visual_sits_at_the_goal_configurations_gripper(built):
    """pandaReaching::SetGoalVisuals (Reaching.cpp:62-85): the target sphere goes where the gripper is at the GOAL joints."""
    tc = make("reaching")
    m = tc.model
    snap, _ = tc.init_snapshot(4)
    joints, bodies, rest = tc.trial_row(4)
    q = np.zeros(m.nq)
    q[tc.qpos_adr[:7]] = rest[:7]
    pos, _ = m.body_pose(q, m.name2id(capi.OBJ_BODY, "franka_gripper"))
    qa = int(m.int("jnt_qposadr")[int(m.int("body_jntadr")[m.name2id(capi.OBJ_BODY, "target")])])
    assert np.array_equal(m.split(snap)["qpos"][qa:qa + 3], pos)


def test_first_control_is_pinv_jacobian_times_gain_error_plus_gravity_compensation(built):
    """One controller step recomputed in numpy from the oracle's own kinematic fields: numpy.linalg.pinv of the 6 x 7
    body Jacobian (an SVD - independent of the oracle's Householder solve), the quat2Axis orientation error, the float
    gains {100, 100, 200, 80, 80, 80} and qfrc_bias (PushBaseClass.cpp:139-248)."""
    tc = make("box_sweep")
    m = tc.model
    om = oracle_model(tc)
    T = 40
    for trial in (0, 5):
        snap, goal = tc.init_snapshot(trial)
        ot = oracle_task(tc, target=tc.load_trial(trial)[1])
        snap0, U, _, warn = oracle.init_sequence_ee(om, ot, snap, goal, 0, T)
        assert warn == 0
        d = oracle.Data(om)
        d.load(snap0)
        d.forward()
        ee = m.name2id(capi.OBJ_BODY, "franka_gripper")
        obj = m.name2id(capi.OBJ_BODY, "goal")
        xpos, xquat = d.get("xpos")[3 * ee:3 * ee + 3], d.get("xquat")[4 * ee:4 * ee + 4]
        cdof, com = d.get("cdof").reshape(-1, 6), d.get("subtree_com").reshape(-1, 3)
        root = int(m.int("body_rootid")[ee])
        J = np.zeros((6, 7))
        for i in range(7):  # the gripper hangs off all seven arm joints
            J[:3, i] = cdof[i, 3:] + np.cross(cdof[i, :3], xpos - com[root])
            J[3:, i] = cdof[i, :3]
        qa = int(m.int("jnt_qposadr")[int(m.int("body_jntadr")[obj])])
        obj_pos = m.split(snap0)["qpos"][qa:qa + 3]
        ang = np.arctan2(goal[1] - obj_pos[1], goal[0] - obj_pos[0]) - REF_PI / 4
        if ang < -(REF_PI / 2):
            ang += 2 * REF_PI
        R = np.array([[np.cos(ang), np.sin(ang), 0.0], [np.sin(ang), -np.cos(ang), 0.0], [0.0, 0.0, -1.0]])  # columns x, y = z cross x, z
        assert abs(np.linalg.det(R) - 1.0) < 1e-12
        # rotation matrix -> quaternion the long way round (eigenvector of the rotation), independent of Eigen's branchy formula
        w_, v_ = np.linalg.eig(R)
        axis = np.real(v_[:, np.argmin(np.abs(w_ - 1.0))])
        angle = np.arccos(np.clip((np.trace(R) - 1) / 2, -1, 1))
        if np.dot(np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]]), axis) < 0:
            axis = -axis
        qd = np.concatenate([[np.cos(angle / 2)], np.sin(angle / 2) * axis])
        if np.dot(qd, xquat) < 0:
            qd = -qd
        qi = xquat * np.array([1, -1, -1, -1])
        w = qd[0] * qi[0] - qd[1:] @ qi[1:]
        v = qd[0] * qi[1:] + qi[0] * qd[1:] + np.cross(qd[1:], qi[1:])
        s = np.sqrt(1 - w * w)
        rot_err = (v if s < 0.001 else v / s) * 2 * np.arccos(min(w, 1.0))
        err = np.concatenate([np.zeros(3), rot_err])  # first path point = the gripper's own position
        u = np.linalg.pinv(J) @ (err * np.array([100, 100, 200, 80, 80, 80.0])) + d.get("qfrc_bias")[:7]
        assert np.abs(u - U[0]).max() < 1e-9 * max(1.0, np.abs(u).max()), (trial, u, U[0])


def test_box_sweep_follower_reaches_the_box_and_ilqr_progresses(built):
    tc = make("box_sweep", kp_method=capi.KP_SET_INTERVAL, min_N=5)
    om = oracle_model(tc)
    T = 300
    snap, goal = tc.init_snapshot(0)
    ot = oracle_task(tc, target=tc.load_trial(0)[1])
    snap0, U, _, warn = oracle.init_sequence_ee(om, ot, snap, goal, 0, T)
    ncon, ee, _ = _play(om, tc, snap0, U)
    assert ncon > 8  # box on floor (4) + display_goal (4) + the gripper on the box
    assert abs(ee[-1, 2] - 0.28) < 0.03  # way-point height 0.28f
    il = oracle.ILQR(om, ot, T, threads=8)
    il.set_lambda(0.01)
    il.optimise(snap0, U, 4, 2)
    assert il.scalars()["cost_reduction"] > 0.2


def test_push_setup_parks_the_gripper_behind_the_cylinder(built):
    """EEWayPointsSetup (PushBaseClass.cpp:8-44): after the setup controls the gripper sits 5 cm behind the cylinder on
    the line to the goal, at z = 0.28; the optimisation controls then move it towards the goal."""
    tc = make("push_ncl")
    m = tc.model
    om = oracle_model(tc)
    snap, goal = tc.init_snapshot(2)
    ot = oracle_task(tc, target=tc.load_trial(2)[1])
    setup, T = 400, 150
    snap0, U, Us, warn = oracle.init_sequence_ee(om, ot, snap, goal, setup, T)
    assert warn == 0 and Us.shape == (setup, 7)
    f = m.split(snap0)
    ee = m.name2id(capi.OBJ_BODY, "franka_gripper")
    obj = m.name2id(capi.OBJ_BODY, "goal")
    qa = int(m.int("jnt_qposadr")[int(m.int("body_jntadr")[obj])])
    pos, _ = m.body_pose(f["qpos"], ee)
    o = f["qpos"][qa:qa + 3]
    ang = np.arctan2(goal[1] - o[1], goal[0] - o[0])
    want = np.array([o[0] - 0.05 * np.cos(ang), o[1] - 0.05 * np.sin(ang), 0.28])
    assert np.abs(pos - want).max() < 0.06, (pos, want)  # a proportional follower lags its moving target by a few cm
    ncon, ee_path, d = _play(om, tc, snap0, U)
    moved = ee_path[-1, :2] - ee_path[0, :2]
    assert np.dot(moved, goal - o[:2]) > 0  # towards the goal


def test_unsupported_pairs_fail_at_load_time(built):
    """pair types without a narrow-phase function used to run silently with the contact missing."""
    xml = """<mujoco><worldbody>
      <body pos="0 0 1"><freejoint/><geom name="a" type="sphere" size="0.1"/></body>
      <body pos="0 0 2"><freejoint/><geom name="b" type="sphere" size="0.1"/></body>
    </worldbody></mujoco>"""
    import ctypes as C
    h = C.c_void_p()
    rc = capi.lib().tokp_model_compile_string(xml.encode(), b".", 0.01, C.byref(h))
    assert rc == 2  # TOKP_ERR_MODEL
    assert b"unsupported collision pair" in capi.lib().tokp_last_error()


def _compile_rc(xml, base=b"."):
    import ctypes as C
    h = C.c_void_p()
    rc = capi.lib().tokp_model_compile_string(xml.encode(), base, 0.01, C.byref(h))
    return rc, capi.lib().tokp_last_error()


def test_dynamics_outside_the_subset_fail_at_load_time(built):
    """frictionloss / joint springs / tendons / equality constraints change the dynamics and are not
    modelled by the step; a model that uses them must not load as if they were absent."""
    body = '<body pos="0 0 1"><joint type="hinge" axis="0 1 0" %s/><geom type="sphere" size="0.1"/></body>'
    ok = "<mujoco><worldbody>%s</worldbody></mujoco>" % (body % 'damping="0.1"')
    assert _compile_rc(ok)[0] == 0
    for attr in ('frictionloss="0.2"', 'stiffness="3"'):
        rc, err = _compile_rc("<mujoco><worldbody>%s</worldbody></mujoco>" % (body % attr))
        assert rc == 2 and b"outside the supported subset" in err, (attr, err)
    # a default class nobody uses may carry the attribute (the Franka scenes' panda_overlay does)
    unused = '<mujoco><default><default class="x"><joint frictionloss="10"/></default></default><worldbody>%s</worldbody></mujoco>'
    assert _compile_rc(unused % (body % ""))[0] == 0
    for section in ('<tendon><fixed name="t"/></tendon>', '<equality><weld body1="a"/></equality>'):
        rc, err = _compile_rc("<mujoco><worldbody>%s</worldbody>%s</mujoco>" % (body % "", section))
        assert rc == 2 and b"outside the supported subset" in err, (section, err)


def test_obj_face_index_outside_the_vertex_list_is_rejected(built, tmp_path):
    """read_obj used the indices unchecked (out-of-bounds read in the inertia pass)."""
    (tmp_path / "bad.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nv 0 0 1\nf 1 2 3\nf 1 2 4\nf 1 3 4\nf 2 3 9\n")
    xml = ('<mujoco><asset><mesh name="m" file="bad.obj"/></asset><worldbody><geom type="plane" size="1 1 1"/>'
           '<body pos="0 0 1"><freejoint/><geom type="mesh" mesh="m"/></body></worldbody></mujoco>')
    rc, err = _compile_rc(xml, str(tmp_path).encode())
    assert rc == 2 and b"face index outside" in err, err
