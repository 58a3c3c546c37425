// Device restatement of the task residuals / cost (MT::Residuals overrides:
// src/ModelTranslator/Acrobot.cpp:26-55, src/ModelTranslator/Walker.cpp:59-90, src/ModelTranslator/Reaching.cpp:28-60;
// MT::CostFunction src/ModelTranslator/ModelTranslator.cpp:314-327).
#pragma once
#include "traj.h"

namespace tokp {

// q, v: full qpos / qvel of the world; u: full ctrl vector
__device__ inline void task_residuals(const TaskDev& t, const double* q, const double* v, const double* u,
                                      const double* target, double* r) {
  switch (t.kind) {
    case TOKP_TASK_ACROBOT:
      r[0] = q[t.qpos_adr[0]] - target[0];
      r[1] = q[t.qpos_adr[1]] - target[1];
      r[2] = v[t.dof_adr[0]] - target[2];
      r[3] = v[t.dof_adr[1]] - -target[3];  // sic (Acrobot.cpp:46)
      r[4] = u[t.ctrl_adr[0]] - target[4];
      break;
    case TOKP_TASK_WALKER:
      r[0] = q[t.qpos_adr[0]] - target[0];
      r[1] = q[t.qpos_adr[2]] - target[1];
      r[2] = v[t.dof_adr[1]] - target[2];
      for (int i = 0; i < t.m; i++) r[3 + i] = u[t.ctrl_adr[i]] - target[3 + i];
      break;
    case TOKP_TASK_REACHING:  // Reaching.cpp:28-60: joint positions against the goal configuration, then velocities
      for (int i = 0; i < t.n; i++) { r[i] = q[t.qpos_adr[i]] - target[i]; r[t.n + i] = v[t.dof_adr[i]]; }
      break;
    default:
      for (int i = 0; i < t.nr; i++) r[i] = 0.0;
  }
}

__device__ inline double task_cost(const TaskDev& t, const double* r, bool terminal) {
  const double* w = terminal ? t.weight_terminal : t.weight;
  double c = 0.0;
  for (int i = 0; i < t.nr; i++) c = __dadd_rn(c, __dmul_rn(w[i], __dmul_rn(r[i], r[i])));
  return c;
}

constexpr int kMaxResid = 16;

}  // namespace tokp
