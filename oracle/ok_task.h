// ORACLE (test infrastructure only).  Task description + ModelTranslator restatement:
// state-vector <-> mjData mapping, residuals, cost (reference: src/ModelTranslator/ModelTranslator.cpp
// :314-327 CostFunction, :552-583 CostDerivativesFromResiduals, :882-1245 Return*Vector, :1466 SetVelocityVector,
// :1548-1610 ComputeStateDofAdrIndices; src/ModelTranslator/Acrobot.cpp:26-55; Walker.cpp:59-90;
// snapshot layout: MuJoCoHelper::CpMjData, src/PhysicsSimulators/MuJoCoHelper.cpp:732-741).
#pragma once
#include <vector>

#include "ok_dyn.h"

namespace ok {

enum { TASK_ACROBOT = 0, TASK_WALKER = 1, TASK_REACHING = 2 };
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
};

struct Task {
  int kind, n, m, num_resid, kp_method, min_N, max_N;
  std::vector<int> qpos_adr, dof_adr, ctrl_adr;
  std::vector<double> target, weight, weight_terminal, jerk_thr, vel_thr, ctrl_limits;
  double iter_thr;

  void init(const Model* mdl, const TaskDesc* d) {
    kind = d->kind; n = d->n; m = d->m; num_resid = d->num_resid;
    kp_method = d->kp_method; min_N = d->min_N; max_N = d->max_N; iter_thr = d->iterative_error_threshold;
    qpos_adr.assign(d->qpos_adr, d->qpos_adr + n);
    dof_adr.assign(d->dof_adr, d->dof_adr + n);
    ctrl_adr.assign(d->ctrl_adr, d->ctrl_adr + m);
    target.assign(d->resid_target, d->resid_target + num_resid);
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
