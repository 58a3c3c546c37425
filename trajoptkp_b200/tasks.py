"""Host-side mirror of ModelTranslator::InitModelTranslator for the Python harness
(reference: src/ModelTranslator/ModelTranslator.cpp:18-96; YAML schema src/FileHandler/FileHandler.cpp:21-289;
trial fixtures src/FileHandler/FileHandler.cpp:471-578).  Produces the arrays behind a tokp_task_desc.
"""
import os

import numpy as np
import yaml

from . import capi

ASSETS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "assets")

# task name -> (task yaml, model xml, task kind)
TASKS = {
    "acrobot": ("tasks/acrobot.yaml", "models/acrobot.xml", capi.TASK_ACROBOT),
    "walker_run": ("tasks/walker_run.yaml", "models/walker_plane.xml", capi.TASK_WALKER),
    "reaching": ("tasks/reaching.yaml", "models/franka/reaching_scene.xml", capi.TASK_REACHING),
    "box_sweep": ("tasks/box_sweep.yaml", "models/franka/boxSweep.xml", capi.TASK_BOXSWEEP),
    "push_ncl": ("tasks/push_ncl.yaml", "models/franka/cylinder_pushing.xml", capi.TASK_PUSH),
    "push_lcl": ("tasks/push_lcl.yaml", "models/franka/cylinder_pushing_lowClutter.xml", capi.TASK_PUSH),
    "push_mcl": ("tasks/push_mcl.yaml", "models/franka/cylinder_pushing_heavyClutter.xml", capi.TASK_PUSH),
}


class TaskConfig:
    """Everything the reference reads from TaskConfigs/<task>.yaml, resolved against a compiled model."""

    def __init__(self, name, model=None, kp_method=None, min_N=None, max_N=None):
        yml, xml, kind = TASKS[name]
        with open(os.path.join(ASSETS, yml)) as f:
            cfg = yaml.safe_load(f)
        self.name, self.kind, self.cfg = name, kind, cfg
        self.timestep = float(cfg.get("timeStep", 0.004))
        self.openloop_horizon = int(cfg.get("openloop_horizon", 3000))
        self.mpc_horizon = int(cfg.get("mpc_horizon", 100))
        self.model_path = os.path.join(ASSETS, xml)
        self.model = model if model is not None else capi.Model(self.model_path, self.timestep)
        m = self.model
        qpos_adr, dof_adr, ctrl_adr, jerk, vel, start = [], [], [], [], [], []
        jnt_qposadr, jnt_dofadr = m.int("jnt_qposadr"), m.int("jnt_dofadr")
        for rname, robot in cfg["robots"].items():
            for j, jn in enumerate(robot["jointNames"]):
                jid = m.name2id(capi.OBJ_JOINT, jn)
                if jid < 0:
                    raise KeyError("joint %s not in model" % jn)
                qpos_adr.append(int(jnt_qposadr[jid]))
                dof_adr.append(int(jnt_dofadr[jid]))
                jerk.append(float(robot["jointJerkThresholds"][j]))
                vel.append(float(robot["magVelThresholds"][j]))
                start.append(float(robot["startPos"][j]))
            for an in robot["actuatorNames"]:
                aid = m.name2id(capi.OBJ_ACTUATOR, an)
                if aid < 0:
                    raise KeyError("actuator %s not in model" % an)
                ctrl_adr.append(aid)
        # rigid bodies (stateVectorList::Update, include/StdInclude.h:119-188; ComputeStateDofAdrIndices,
        # ModelTranslator.cpp:1548-1610): per body the active linear x, y, z then the active angular DoFs of its free
        # joint.  Linear elements read qpos directly; angular ones would need the axis-angle read-out (quat2Axis) and
        # are rejected until a shipped task in scope activates them.
        self.num_robot_dofs = len(qpos_adr)
        self.body_names, self.body_start = [], []
        body_jntadr = m.int("body_jntadr")
        for bname, body in (cfg.get("bodies") or {}).items():
            bid = m.name2id(capi.OBJ_BODY, bname)
            if bid < 0:
                raise KeyError("body %s not in model" % bname)
            if any(body["activeAngularDOF"]):
                raise NotImplementedError("angular rigid-body state elements are not supported yet")
            jid = int(body_jntadr[bid])
            self.body_names.append(bname)
            self.body_start.append((int(jnt_qposadr[jid]), list(map(float, body["startLinearPos"])),
                                    list(map(float, body["startAngularPos"]))))
            for k in range(3):
                if body["activeLinearDOF"][k]:
                    qpos_adr.append(int(jnt_qposadr[jid]) + k)
                    dof_adr.append(int(jnt_dofadr[jid]) + k)
                    jerk.append(float(body["linearJerkThreshold"][k]))
                    vel.append(float(body["linearMagVelThreshold"][k]))
                    start.append(float(body["startLinearPos"][k]))
        self.aux = np.zeros(0, dtype=np.int32)
        if kind in (capi.TASK_BOXSWEEP, capi.TASK_PUSH):
            self.aux = np.array([m.name2id(capi.OBJ_BODY, "goal"), m.name2id(capi.OBJ_BODY, "franka_gripper")],
                                dtype=np.int32)
            if (self.aux < 0).any():
                raise KeyError("box_sweep / push need the bodies 'goal' and 'franka_gripper'")
        target, weight, weight_t = [], [], []
        self.resid_target_sizes = []
        for rn, r in cfg["residuals"].items():
            dim = int(r.get("resid_dimension", 1))
            if kind in (capi.TASK_BOXSWEEP, capi.TASK_PUSH):  # vector targets, concatenated (tokp_task_desc.num_target)
                target += [float(x) for x in r["target"]]
                weight.append(float(r["weight"]))
                weight_t.append(float(r["weight_terminal"]))
                self.resid_target_sizes.append(len(r["target"]))
                continue
            for k in range(dim):  # FileHandler.cpp:276-287 replicates the entry resid_dimension times
                # pandaReaching::Residuals reads residual_list[0].target[i] for its i-th residual (Reaching.cpp:46)
                first_group = kind == capi.TASK_REACHING and len(target) < dim
                target.append(float(r["target"][k if first_group else 0]))
                weight.append(float(r["weight"]))
                weight_t.append(float(r["weight_terminal"]))
                self.resid_target_sizes.append(len(r["target"]))
        self.qpos_adr = np.array(qpos_adr, dtype=np.int32)
        self.dof_adr = np.array(dof_adr, dtype=np.int32)
        self.ctrl_adr = np.array(ctrl_adr, dtype=np.int32)
        self.target = np.array(target)
        self.weight = np.array(weight)
        self.weight_terminal = np.array(weight_t)
        self.jerk_thr = np.array(jerk)
        self.vel_thr = np.array(vel)
        self.start_pos = np.array(start)
        self.n, self.m, self.nr = len(qpos_adr), len(ctrl_adr), len(weight)
        self.kp_method = capi.KP_NAMES[cfg.get("keypointMethod", "set_interval")] if kp_method is None else kp_method
        self.min_N = int(cfg.get("minN", 1)) if min_N is None else min_N
        self.max_N = int(cfg.get("maxN", 10)) if max_N is None else max_N
        self.iter_thr = float(cfg.get("iterativeErrorThreshold", 10))

    def desc(self):
        d = capi.TaskDesc()
        d.kind, d.n, d.m = self.kind, self.n, self.m
        d.qpos_adr, d.dof_adr, d.ctrl_adr = capi._i(self.qpos_adr), capi._i(self.dof_adr), capi._i(self.ctrl_adr)
        d.num_resid = self.nr
        d.resid_target, d.resid_weight = capi._d(self.target), capi._d(self.weight)
        d.resid_weight_terminal = capi._d(self.weight_terminal)
        d.kp_method, d.min_N, d.max_N = self.kp_method, self.min_N, self.max_N
        d.jerk_thresholds, d.velchange_thresholds = capi._d(self.jerk_thr), capi._d(self.vel_thr)
        d.iterative_error_threshold = self.iter_thr
        d.num_target = len(self.target) if len(self.target) != self.nr else 0
        d.num_aux = len(self.aux)
        d.aux = capi._i(self.aux) if len(self.aux) else None
        return d

    # ---- TestTasks fixtures --------------------------------------------------------------------
    def num_trials(self):
        path = os.path.join(ASSETS, "test_tasks", self.name + ".csv")
        return sum(1 for l in open(path) if l.strip() and not l.startswith("#"))

    def load_trial(self, i):
        """Row i of TestTasks/<task>: (start joint positions, residual targets[num_resid])."""
        path = os.path.join(ASSETS, "test_tasks", self.name + ".csv")
        rows = [l.strip() for l in open(path) if l.strip() and not l.startswith("#")]
        vals = [float(v) for v in rows[i].split(",")]
        # the file lists every robot joint, then x, y, z, roll, pitch, yaw of every body (FileHandler.cpp:536-553)
        nfile = self.num_robot_dofs + 6 * len(self.body_names)
        start = np.array(vals[:self.num_robot_dofs])
        for k in range(len(self.body_names)):
            o = self.num_robot_dofs + 6 * k
            if any(abs(a) > 0 for a in vals[o + 3:o + 6]):
                raise NotImplementedError("rotated rigid-body start poses are not supported yet")
            start = np.concatenate([start, vals[o:o + 3]])
        rest = vals[nfile:]
        if self.kind == capi.TASK_BOXSWEEP:
            return start, np.array(rest[:len(self.target)])
        if self.kind == capi.TASK_PUSH:
            tg = np.array(rest[:len(self.target)])
            # obstacle residuals measure the displacement from the obstacle's START position
            # (full_state_vector.rigid_bodies[i + 1].start_linear_pos, TwoDPushing.cpp:332-333), not a file target
            for i in range(self.nr - 4):
                tg[4 + 2 * i:6 + 2 * i] = start[10 + 3 * i:12 + 3 * i]
            return start, tg
        targets, p = [], 0
        for k, sz in enumerate(self.resid_target_sizes):
            if self.kind == capi.TASK_REACHING:  # residual_list[0].target[k] for the position residuals, 0 after
                targets.append(rest[k] if k < self.n else 0.0)
            else:
                targets.append(rest[p])  # residual_list[k].target[0]
            p += sz
        return start, np.array(targets)

    # ---- ModelTranslator::InitialiseSystemToStartState ------------------------------------------
    SETUP_HORIZON = 1000  # GenTestingData.cpp:197: CreateInitSetupControls(1000)

    def trial_row(self, i):
        """Raw row i of TestTasks/<task>: (robot joint starts, per body [x, y, z, r, p, y], residual targets as filed)."""
        path = os.path.join(ASSETS, "test_tasks", self.name + ".csv")
        rows = [l.strip() for l in open(path) if l.strip() and not l.startswith("#")]
        vals = [float(v) for v in rows[i].rstrip(",").split(",")]
        nr = self.num_robot_dofs
        bodies = [vals[nr + 6 * k:nr + 6 * k + 6] for k in range(len(self.body_names))]
        return np.array(vals[:nr]), bodies, vals[nr + 6 * len(self.body_names):]

    def init_snapshot(self, i):
        """The mjData state ModelTranslator::InitialiseSystemToStartState (ModelTranslator.cpp:1711-1781) leaves for
        trial i of TestTasks (FileHandler::LoadTaskFromFile, FileHandler.cpp:471-578): everything zero, then
        SetGoalVisuals of the task, robot joints at their start positions, every rigid body of the state vector at its
        start pose (axis2Quat of the filed angles, StdInclude.cpp:124-147).  Returns (snapshot, goal_xy)."""
        m = self.model
        joints, bodies, rest = self.trial_row(i)
        qpos = np.zeros(m.nq)
        jnt_qposadr, body_jntadr = m.int("jnt_qposadr"), m.int("body_jntadr")

        def set_body(name, pos, ang):
            bid = m.name2id(capi.OBJ_BODY, name)
            if bid < 0:
                raise KeyError("body %s not in model" % name)
            qa = int(jnt_qposadr[int(body_jntadr[bid])])
            angle = float(np.sqrt(ang[0] ** 2 + ang[1] ** 2 + ang[2] ** 2))
            sh = np.sin(angle / 2.0)
            qpos[qa:qa + 3] = pos
            qpos[qa + 3:qa + 7] = [np.cos(angle / 2.0), ang[0] * sh, ang[1] * sh, ang[2] * sh]

        goal_xy = np.array(rest[:2]) if self.kind in (capi.TASK_BOXSWEEP, capi.TASK_PUSH) else None
        # SetGoalVisuals (BoxSweep.cpp:52-61: z = 0; TwoDPushing.cpp:213-223: z = 0.032; Reaching.cpp:62-85)
        if self.kind == capi.TASK_BOXSWEEP:
            set_body("display_goal", [goal_xy[0], goal_xy[1], 0.0], [0.0, 0.0, 0.0])
        elif self.kind == capi.TASK_PUSH:
            set_body("display_goal", [goal_xy[0], goal_xy[1], 0.032], [0.0, 0.0, 0.0])
        elif self.kind == capi.TASK_REACHING:
            q = qpos.copy()
            q[self.qpos_adr[:7]] = rest[:7]
            pos, _ = m.body_pose(q, m.name2id(capi.OBJ_BODY, "franka_gripper"))
            set_body("target", pos, [0.0, 0.0, 0.0])
        qpos[self.qpos_adr[:self.num_robot_dofs]] = joints
        for name, b in zip(self.body_names, bodies):
            set_body(name, b[:3], b[3:])
        return m.snapshot(qpos=qpos), goal_xy

    def start_snapshot(self, start_pos=None, qvel=None, ctrl=None):
        """InitialiseSystemToStartState for joint-only state vectors: qpos <- start positions, everything else 0."""
        m = self.model
        qpos = m.real("qpos0").copy()
        qpos[self.qpos_adr] = self.start_pos if start_pos is None else start_pos
        v = np.zeros(m.nv)
        if qvel is not None:
            v[self.dof_adr] = qvel
        return m.snapshot(qpos=qpos, qvel=v, ctrl=ctrl)
