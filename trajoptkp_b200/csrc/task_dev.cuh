// Device restatement of the task residuals / cost (MT::Residuals overrides:
// src/ModelTranslator/Acrobot.cpp:26-55, src/ModelTranslator/Walker.cpp:59-90, src/ModelTranslator/Reaching.cpp:28-60, src/ModelTranslator/BoxSweep.cpp:101-139, src/ModelTranslator/TwoDPushing.cpp:291-356;
// MT::CostFunction src/ModelTranslator/ModelTranslator.cpp:314-327).
#pragma once
#include "step.cuh"
#include "traj.h"

namespace tokp {

// World position of one body from qpos alone (mj_kinematics restricted to the chain world -> body; the reference's
// residuals call mj_kinematics and read xpos: BoxSweep.cpp:104-126, MuJoCoHelper::GetBodyPoseQuatViaXpos).
__device__ inline void body_world_pos(const DM& m, const double* qpos, int body, double* out) {
  int chain[32], nc = 0;
  for (int b = body; b > 0 && nc < 32; b = m.body_parentid[b]) chain[nc++] = b;
  double xpos[3] = {0.0, 0.0, 0.0}, xquat[4] = {1.0, 0.0, 0.0, 0.0};
  for (int c = nc - 1; c >= 0; c--) {
    int i = chain[c];
    int jntadr = m.body_jntadr[i], jntnum = m.body_jntnum[i];
    if (jntnum == 1 && m.jnt_type[jntadr] == tk::JNT_FREE) {
      int qa = m.jnt_qposadr[jntadr];
      for (int k = 0; k < 3; k++) xpos[k] = qpos[qa + k];
      for (int k = 0; k < 4; k++) xquat[k] = qpos[qa + 3 + k];
      tk::normalize4(xquat);
      continue;
    }
    double v[3], q[4];
    tk::rot_vec_quat(v, &m.body_pos[3 * i], xquat);
    for (int k = 0; k < 3; k++) xpos[k] += v[k];
    tk::mul_quat(q, xquat, &m.body_quat[4 * i]);
    for (int k = 0; k < 4; k++) xquat[k] = q[k];
    for (int j = jntadr; j < jntadr + jntnum; j++) {
      int qa = m.jnt_qposadr[j];
      double xaxis[3], xanchor[3];
      tk::rot_vec_quat(xaxis, &m.jnt_axis[3 * j], xquat);
      tk::rot_vec_quat(xanchor, &m.jnt_pos[3 * j], xquat);
      for (int k = 0; k < 3; k++) xanchor[k] += xpos[k];
      if (m.jnt_type[j] == tk::JNT_SLIDE) {
        double dq = qpos[qa] - m.qpos0[qa];
        for (int k = 0; k < 3; k++) xpos[k] += xaxis[k] * dq;
      } else if (m.jnt_type[j] == tk::JNT_HINGE) {
        double qloc[4], tmp[4], vec[3];
        double jaxis[3] = {m.jnt_axis[3 * j], m.jnt_axis[3 * j + 1], m.jnt_axis[3 * j + 2]};
        tk::axis_angle2quat(qloc, jaxis, qpos[qa] - m.qpos0[qa]);
        tk::mul_quat(tmp, xquat, qloc);
        for (int k = 0; k < 4; k++) xquat[k] = tmp[k];
        tk::rot_vec_quat(vec, &m.jnt_pos[3 * j], xquat);
        for (int k = 0; k < 3; k++) xpos[k] = xanchor[k] - vec[k];
      }
    }
    tk::normalize4(xquat);
  }
  for (int k = 0; k < 3; k++) out[k] = xpos[k];
}

// q, v: full qpos / qvel of the world; u: full ctrl vector
__device__ inline void task_residuals(const DM& dm, const TaskDev& t, const double* q, const double* v, const double* u,
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
    case TOKP_TASK_BOXSWEEP: {
      // BoxSweep.cpp:101-139: state = 7 arm joints, then the goal body's linear x, y, z (free joint); targets =
      // goal_pos (x, y), goal_vel (x, y), reach (d)
      int gq = t.qpos_adr[t.n - 3], gv = t.dof_adr[t.n - 3];
      double dx = q[gq] - target[0], dy = q[gq + 1] - target[1];
      r[0] = sqrt(dx * dx + dy * dy);
      dx = v[gv] - target[2]; dy = v[gv + 1] - target[3];
      r[1] = sqrt(dx * dx + dy * dy);
      double ee[3];
      body_world_pos(dm, q, t.aux[1], ee);
      dx = ee[0] - q[gq]; dy = ee[1] - q[gq + 1];
      double dz = ee[2] - q[gq + 2];
      r[2] = sqrt(dx * dx + dy * dy + dz * dz) - target[4];
      break;
    }
    case TOKP_TASK_PUSH: {
      // TwoDPushing.cpp:291-356: state = 7 arm joints, then x, y, z of the goal and of every obstacle (free joints)
      int nobs = t.nr - 4;
      int gq = t.qpos_adr[7], gv = t.dof_adr[7];
      double dx = q[gq] - target[0], dy = q[gq + 1] - target[1];
      int k = 0;
      r[k++] = sqrt(dx * dx + dy * dy);
      r[k++] = sqrt(v[gv] * v[gv] + v[gv + 1] * v[gv + 1]);
      for (int i = 0; i < nobs; i++) {  // obstacle displacement from its start position
        int oq = t.qpos_adr[10 + 3 * i];
        dx = q[oq] - target[4 + 2 * i]; dy = q[oq + 1] - target[5 + 2 * i];
        r[k++] = sqrt(dx * dx + dy * dy);
      }
      r[k++] = v[t.dof_adr[5]];
      double ee[3];
      body_world_pos(dm, q, t.aux[1], ee);
      dx = ee[0] - q[gq]; dy = ee[1] - q[gq + 1];
      double dz = ee[2] - q[gq + 2];
      r[k++] = sqrt(dx * dx + dy * dy + dz * dz);
      break;
    }
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
