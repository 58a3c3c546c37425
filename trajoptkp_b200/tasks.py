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
        if cfg.get("bodies"):
            raise NotImplementedError("rigid-body state elements are outside round-1 scope")
        target, weight, weight_t = [], [], []
        self.resid_target_sizes = []
        for rn, r in cfg["residuals"].items():
            dim = int(r.get("resid_dimension", 1))
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
        self.n, self.m, self.nr = len(qpos_adr), len(ctrl_adr), len(target)
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
        start = np.array(vals[:self.n])
        rest = vals[self.n:]
        targets, p = [], 0
        for k, sz in enumerate(self.resid_target_sizes):
            if self.kind == capi.TASK_REACHING:  # residual_list[0].target[k] for the position residuals, 0 after
                targets.append(rest[k] if k < self.n else 0.0)
            else:
                targets.append(rest[p])  # residual_list[k].target[0]
            p += sz
        return start, np.array(targets)

    def start_snapshot(self, start_pos=None, qvel=None, ctrl=None):
        """InitialiseSystemToStartState for joint-only state vectors: qpos <- start positions, everything else 0."""
        m = self.model
        qpos = m.real("qpos0").copy()
        qpos[self.qpos_adr] = self.start_pos if start_pos is None else start_pos
        v = np.zeros(m.nv)
        if qvel is not None:
            v[self.dof_adr] = qvel
        return m.snapshot(qpos=qpos, qvel=v, ctrl=ctrl)
