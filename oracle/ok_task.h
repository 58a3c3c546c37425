// ORACLE (test infrastructure only).  Task description + ModelTranslator restatement:
// state-vector <-> mjData mapping, residuals, cost (reference: src/ModelTranslator/ModelTranslator.cpp
// :314-327 CostFunction, :552-583 CostDerivativesFromResiduals, :882-1245 Return*Vector, :1466 SetVelocityVector,
// :1548-1610 ComputeStateDofAdrIndices; src/ModelTranslator/Acrobot.cpp:26-55; Walker.cpp:59-90;
// snapshot layout: MuJoCoHelper::CpMjData, src/PhysicsSimulators/MuJoCoHelper.cpp:732-741).
#pragma once
#include <vector>

#include "ok_dyn.h"

namespace ok {

enum { TASK_ACROBOT = 0, TASK_WALKER = 1, TASK_REACHING = 2, TASK_BOXSWEEP = 3, TASK_PUSH = 4 };
enum { KP_SET_INTERVAL = 0, KP_ADAPTIVE_JERK = 1, KP_ADAPTIVE_ACCEL = 2, KP_ITERATIVE_ERROR = 3, KP_VELOCITY_CHANGE = 4 };

// Mirrors tokp_task_desc (include/tokp.h) field for field.
struct TaskDesc {
  int kind, n, m;
  const int* qpos_adr;   // [n]
  const int* dof_adr;    // [n]  (what the reference calls StateIndexToQposIndex)
  const int* ctrl_adr;   // [m]
  int num_resid;
  const double* resid_target;           // [num_resid]
  const double* resid_weight;           // [num_resid]
  const double* resid_weight_terminal;  // [num_resid]
  int kp_method, min_N, max_N;
  const double* jerk_thresholds;        // [n]
  const double* velchange_thresholds;   // [n]
  double iterative_error_threshold;
  int num_target;  // 0 = num_resid; else resid_target holds the residuals' target vectors concatenated
  int num_aux;
  const int* aux;  // BoxSweep: {goal body id, franka_gripper body id}
};

struct Task {
  int kind, n, m, num_resid, kp_method, min_N, max_N;
  std::vector<int> qpos_adr, dof_adr, ctrl_adr;
  std::vector<double> target, weight, weight_terminal, jerk_thr, vel_thr, ctrl_limits;
  double iter_thr;
  std::vector<int> aux;

  void init(const Model* mdl, const TaskDesc* d) {
    int ntg = d->num_target > 0 ? d->num_target : d->num_resid;
    if (d->aux) aux.assign(d->aux, d->aux + d->num_aux);
    kind = d->kind; n = d->n; m = d->m; num_resid = d->num_resid;
    kp_method = d->kp_method; min_N = d->min_N; max_N = d->max_N; iter_thr = d->iterative_error_threshold;
    qpos_adr.assign(d->qpos_adr, d->qpos_adr + n);
    dof_adr.assign(d->dof_adr, d->dof_adr + n);
    ctrl_adr.assign(d->ctrl_adr, d->ctrl_adr + m);
    target.assign(d->resid_target, d->resid_target + ntg);
    weight.assign(d->resid_weight, d->resid_weight + num_resid);
    weight_terminal.assign(d->resid_weight_terminal, d->resid_weight_terminal + num_resid);
    jerk_thr.assign(d->jerk_thresholds, d->jerk_thresholds + n);
    vel_thr.assign(d->velchange_thresholds, d->velchange_thresholds + n);
    // ReturnControlLimits: the first 2m entries of model->actuator_ctrlrange (MuJoCoHelper.cpp:262-264)
    ctrl_limits.assign(mdl->actuator_ctrlrange, mdl->actuator_ctrlrange + 2 * m);
  }
};

// flat snapshot: [time | qpos | qvel | qacc | qacc_warmstart | qfrc_applied | xfrc_applied | ctrl]
inline int snapshot_size(const Model* m) { return 1 + m->nq + 4 * m->nv + 6 * m->nbody + m->nu; }

template <class R>
void load_snapshot(const Model* m, Data<R>* d, const double* s) {
  int nq = m->nq, nv = m->nv;
  d->time = R(s[0]);
  const double* p = s + 1;
  for (int i = 0; i < nq; i++) d->qpos[i] = R(p[i]);
  p += nq;
  for (int i = 0; i < nv; i++) d->qvel[i] = R(p[i]);
  p += nv;
  for (int i = 0; i < nv; i++) d->qacc[i] = R(p[i]);
  p += nv;
  for (int i = 0; i < nv; i++) d->qacc_warmstart[i] = R(p[i]);
  p += nv;
  p += nv + 6 * m->nbody;  // applied forces: always zero in scope
  for (int i = 0; i < m->nu; i++) d->ctrl[i] = R(p[i]);
}

template <class R>
void save_snapshot(const Model* m, const Data<R>* d, double* s) {
  int nq = m->nq, nv = m->nv;
  s[0] = to_double(d->time);
  double* p = s + 1;
  for (int i = 0; i < nq; i++) p[i] = to_double(d->qpos[i]);
  p += nq;
  for (int i = 0; i < nv; i++) p[i] = to_double(d->qvel[i]);
  p += nv;
  for (int i = 0; i < nv; i++) p[i] = to_double(d->qacc[i]);
  p += nv;
  for (int i = 0; i < nv; i++) p[i] = to_double(d->qacc_warmstart[i]);
  p += nv;
  for (int i = 0; i < nv + 6 * m->nbody; i++) p[i] = 0.0;
  p += nv + 6 * m->nbody;
  for (int i = 0; i < m->nu; i++) p[i] = to_double(d->ctrl[i]);
}

// ReturnStateVector: [positions(n) ; velocities(n)]
template <class R>
void return_state_vector(const Task& t, const Data<R>* d, R* x) {
  for (int i = 0; i < t.n; i++) { x[i] = d->qpos[t.qpos_adr[i]]; x[t.n + i] = d->qvel[t.dof_adr[i]]; }
}

// xpos of one body from qpos alone: mj_kinematics restricted to the chain world -> body (the reference's residuals
// call mj_kinematics and read d->xpos: BoxSweep.cpp:104-126)
template <class R>
void body_world_pos(const Model* m, const R* qpos, int body, R* out) {
  std::vector<int> chain;
  for (int b = body; b > 0; b = m->body_parentid[b]) chain.push_back(b);
  R xpos[3] = {R(0.0), R(0.0), R(0.0)}, xquat[4] = {R(1.0), R(0.0), R(0.0), R(0.0)};
  for (int c = (int)chain.size() - 1; c >= 0; c--) {
    int i = chain[c];
    int jntadr = m->body_jntadr[i], jntnum = m->body_jntnum[i];
    if (jntnum == 1 && m->jnt_type[jntadr] == JNT_FREE) {
      int qa = m->jnt_qposadr[jntadr];
      for (int k = 0; k < 3; k++) xpos[k] = qpos[qa + k];
      for (int k = 0; k < 4; k++) xquat[k] = qpos[qa + 3 + k];
      normalize4(xquat);
      continue;
    }
    R v[3], q[4], bp[3] = {R(m->body_pos[3 * i]), R(m->body_pos[3 * i + 1]), R(m->body_pos[3 * i + 2])};
    R bq[4] = {R(m->body_quat[4 * i]), R(m->body_quat[4 * i + 1]), R(m->body_quat[4 * i + 2]), R(m->body_quat[4 * i + 3])};
    rot_vec_quat(v, bp, xquat);
    for (int k = 0; k < 3; k++) xpos[k] += v[k];
    mul_quat(q, xquat, bq);
    for (int k = 0; k < 4; k++) xquat[k] = q[k];
    for (int j = jntadr; j < jntadr + jntnum; j++) {
      int qa = m->jnt_qposadr[j];
      R jaxis[3] = {R(m->jnt_axis[3 * j]), R(m->jnt_axis[3 * j + 1]), R(m->jnt_axis[3 * j + 2])};
      R jpos[3] = {R(m->jnt_pos[3 * j]), R(m->jnt_pos[3 * j + 1]), R(m->jnt_pos[3 * j + 2])};
      R xaxis[3], xanchor[3];
      rot_vec_quat(xaxis, jaxis, xquat);
      rot_vec_quat(xanchor, jpos, xquat);
      for (int k = 0; k < 3; k++) xanchor[k] += xpos[k];
      if (m->jnt_type[j] == JNT_SLIDE) {
        R dq = qpos[qa] - R(m->qpos0[qa]);
        for (int k = 0; k < 3; k++) xpos[k] += xaxis[k] * dq;
      } else if (m->jnt_type[j] == JNT_HINGE) {
        R qloc[4], tmp[4], vec[3];
        axis_angle2quat(qloc, jaxis, qpos[qa] - R(m->qpos0[qa]));
        mul_quat(tmp, xquat, qloc);
        for (int k = 0; k < 4; k++) xquat[k] = tmp[k];
        rot_vec_quat(vec, jpos, xquat);
        for (int k = 0; k < 3; k++) xpos[k] = xanchor[k] - vec[k];
      }
    }
    normalize4(xquat);
  }
  for (int k = 0; k < 3; k++) out[k] = xpos[k];
}

// MT::Residuals for the task kinds in scope
template <class R>
void residuals(const Task& t, const Data<R>* d, R* r) {
  auto q = [&](int i) { return d->qpos[t.qpos_adr[i]]; };
  auto v = [&](int i) { return d->qvel[t.dof_adr[i]]; };
  auto u = [&](int i) { return d->ctrl[t.ctrl_adr[i]]; };
  switch (t.kind) {
    case TASK_ACROBOT:
      r[0] = q(0) - R(t.target[0]);
      r[1] = q(1) - R(t.target[1]);
      r[2] = v(0) - R(t.target[2]);
      r[3] = v(1) - -R(t.target[3]);  // sic: Acrobot.cpp:46 has "- -"
      r[4] = u(0) - R(t.target[4]);
      break;
    case TASK_WALKER:
      r[0] = q(0) - R(t.target[0]);
      r[1] = q(2) - R(t.target[1]);
      r[2] = v(1) - R(t.target[2]);
      for (int i = 0; i < t.m; i++) r[3 + i] = u(i) - R(t.target[3 + i]);
      break;
    case TASK_BOXSWEEP: {  // BoxSweep.cpp:101-139
      using std::sqrt;
      int gq = t.qpos_adr[t.n - 3], gv = t.dof_adr[t.n - 3];
      R dx = d->qpos[gq] - R(t.target[0]), dy = d->qpos[gq + 1] - R(t.target[1]);
      r[0] = sqrt(dx * dx + dy * dy);
      dx = d->qvel[gv] - R(t.target[2]); dy = d->qvel[gv + 1] - R(t.target[3]);
      r[1] = sqrt(dx * dx + dy * dy);
      R ee[3];
      body_world_pos(d->m, d->qpos.data(), t.aux[1], ee);
      dx = ee[0] - d->qpos[gq]; dy = ee[1] - d->qpos[gq + 1];
      R dz = ee[2] - d->qpos[gq + 2];
      r[2] = sqrt(dx * dx + dy * dy + dz * dz) - R(t.target[4]);
      break;
    }
    case TASK_PUSH: {  // TwoDPushing.cpp:291-356
      using std::sqrt;
      int nobs = t.num_resid - 4, k = 0;
      int gq = t.qpos_adr[7], gv = t.dof_adr[7];
      R dx = d->qpos[gq] - R(t.target[0]), dy = d->qpos[gq + 1] - R(t.target[1]);
      r[k++] = sqrt(dx * dx + dy * dy);
      r[k++] = sqrt(d->qvel[gv] * d->qvel[gv] + d->qvel[gv + 1] * d->qvel[gv + 1]);
      for (int i = 0; i < nobs; i++) {
        int oq = t.qpos_adr[10 + 3 * i];
        dx = d->qpos[oq] - R(t.target[4 + 2 * i]); dy = d->qpos[oq + 1] - R(t.target[5 + 2 * i]);
        r[k++] = sqrt(dx * dx + dy * dy);
      }
      r[k++] = v(5);
      R ee[3];
      body_world_pos(d->m, d->qpos.data(), t.aux[1], ee);
      dx = ee[0] - d->qpos[gq]; dy = ee[1] - d->qpos[gq + 1];
      R dz = ee[2] - d->qpos[gq + 2];
      r[k++] = sqrt(dx * dx + dy * dy + dz * dz);
      break;
    }
    case TASK_REACHING:  // Reaching.cpp:28-60: joint positions against the goal configuration, then joint velocities
      for (int i = 0; i < t.n; i++) { r[i] = q(i) - R(t.target[i]); r[t.n + i] = v(i); }
      break;
    default: break;
  }
}

// MT::CostFunction
inline double cost_function(const Task& t, const double* r, bool terminal) {
  double cost = 0.0;
  for (int i = 0; i < t.num_resid; i++) cost += (terminal ? t.weight_terminal[i] : t.weight[i]) * (r[i] * r[i]);
  return cost;
}

}  // namespace ok
