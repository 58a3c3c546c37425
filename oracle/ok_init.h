// ORACLE (test infrastructure only - never on the product path).
// Initial-control generators of the pushing tasks, restated literally:
//   PushBaseClass::EEWayPointsSetup / EEWayPointsPush / CreateAllEETransitPoints / JacobianEEControl
//     (src/ModelTranslator/PushBaseClass.cpp:8-248),
//   their callers BoxSweep::CreateInitSetupControls / CreateInitOptimisationControls (src/ModelTranslator/BoxSweep.cpp:63-99)
//   and TwoDPushing::CreateInitSetupControls / CreateInitOptimisationControls (src/ModelTranslator/TwoDPushing.cpp:225-290),
//   MuJoCoHelper::GetJacobian (src/PhysicsSimulators/MuJoCoHelper.cpp:562-587, mj_jacBody),
//   quat2Axis / multQuat / invQuat / rotMat2Quat (src/StdInclude/StdInclude.cpp:91-122,176-233),
//   in the call order of GenTestingData::GenDataOpenloopOptimisation (src/GenTestingData.cpp:177-220).
// The reference works on two mjData objects (main_data, master_reset_data) and copies only the 8 snapshot fields
// between them (MuJoCoHelper::CpMjData), so xpos / xquat / cdof / subtree_com read by the controller are the ones of
// the LAST kinematics evaluation on main_data - one step stale from the second path point on.  The Data objects below
// are used the same way, which reproduces that.
// Eigen's completeOrthogonalDecomposition().pseudoInverse() of the 6 x 7 Jacobian is replaced by the minimum-norm
// solution through a Householder QR of J' (the same pseudo-inverse for a full-rank J; a rank-deficient J raises
// warning bit 64 instead of truncating).  PARITY UNPINNED against the reference's bits (no Eigen / MuJoCo here).
#pragma once
#include <array>
#include <cmath>
#include <vector>

#include "ok_task.h"

namespace ok {

constexpr double kRefPI = 3.14152;  // #define PI 3.14152 (include/StdInclude.h:13)

typedef std::array<double, 3> Pt3;

// CpMjData (MuJoCoHelper.cpp:732-741): the 8 snapshot fields, nothing else
inline void cp_mjdata(Data<double>* dst, const Data<double>* src) {
  dst->time = src->time;
  dst->qpos = src->qpos; dst->qvel = src->qvel; dst->qacc = src->qacc;
  dst->qacc_warmstart = src->qacc_warmstart; dst->ctrl = src->ctrl;
}

// quat2Axis (StdInclude.cpp:91-122)
inline void quat2axis_ref(const double q[4], double out[3]) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  if (w > 1) w = 1;
  double angle = 2 * std::acos(w);
  double s = std::sqrt(1 - (w * w));
  if (s < 0.001) { out[0] = x; out[1] = y; out[2] = z; }
  else { out[0] = x / s; out[1] = y / s; out[2] = z / s; }
  for (int k = 0; k < 3; k++) out[k] *= angle;
}

// multQuat (StdInclude.cpp:200-227)
inline void mult_quat_ref(const double l[4], const double r[4], double out[4]) {
  double wr = l[0], xr = l[1], yr = l[2], zr = l[3], ws = r[0], xs = r[1], ys = r[2], zs = r[3];
  out[0] = (wr * ws) - (xr * xs) - (yr * ys) - (zr * zs);
  out[1] = (ws * xr) + (wr * xs) + (yr * zs) - (zr * ys);
  out[2] = (ws * yr) + (wr * ys) + (zr * xs) - (xr * zs);
  out[3] = (ws * zr) + (wr * zs) + (xr * ys) - (yr * xs);
}

// Eigen::Quaterniond(Matrix3d) as rotMat2Quat uses it (StdInclude.cpp:176-187); R row-major
inline void rotmat2quat_ref(const double R[9], double q[4]) {
  double t = R[0] + R[4] + R[8];
  if (t > 0.0) {
    t = std::sqrt(t + 1.0);
    q[0] = 0.5 * t;
    t = 0.5 / t;
    q[1] = (R[7] - R[5]) * t;
    q[2] = (R[2] - R[6]) * t;
    q[3] = (R[3] - R[1]) * t;
  } else {
    int i = 0;
    if (R[4] > R[0]) i = 1;
    if (R[8] > R[4 * i]) i = 2;
    int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
    q[1 + i] = 0.5 * t;
    t = 0.5 / t;
    q[0] = (R[3 * k + j] - R[3 * j + k]) * t;
    q[1 + j] = (R[3 * j + i] + R[3 * i + j]) * t;
    q[1 + k] = (R[3 * k + i] + R[3 * i + k]) * t;
  }
}

// u = pinv(J) f for a 6 x 7 J: Householder QR of A = J' (7 x 6), R' y = f, u = Q [y; 0].  Returns false when a
// diagonal of R is negligible (rank-deficient Jacobian).
inline bool min_norm_solve_6x7(const double J[6][7], const double f[6], double u[7]) {
  double A[7][6], beta[6], Rd[6];
  for (int r = 0; r < 7; r++)
    for (int c = 0; c < 6; c++) A[r][c] = J[c][r];
  for (int k = 0; k < 6; k++) {
    double nrm = 0.0;
    for (int r = k; r < 7; r++) nrm += A[r][k] * A[r][k];
    nrm = std::sqrt(nrm);
    double alpha = A[k][k] > 0.0 ? -nrm : nrm;
    double v0 = A[k][k] - alpha;
    // v = [v0, A[k+1..][k]]; beta = 2 / (v'v)
    double vv = v0 * v0;
    for (int r = k + 1; r < 7; r++) vv += A[r][k] * A[r][k];
    beta[k] = vv > 0.0 ? 2.0 / vv : 0.0;
    A[k][k] = v0;
    Rd[k] = alpha;
    for (int c = k + 1; c < 6; c++) {
      double s = 0.0;
      for (int r = k; r < 7; r++) s += A[r][k] * A[r][c];
      s *= beta[k];
      for (int r = k; r < 7; r++) A[r][c] -= s * A[r][k];
    }
  }
  double rmax = 0.0;
  for (int k = 0; k < 6; k++) rmax = std::fmax(rmax, std::fabs(Rd[k]));
  bool ok = true;
  for (int k = 0; k < 6; k++)
    if (!(std::fabs(Rd[k]) > 6.0 * 2.220446049250313e-16 * rmax)) ok = false;
  // R' y = f : R(k, c) for c > k sits in A[k][c], the diagonal in Rd
  double y[7];
  for (int k = 0; k < 6; k++) {
    double s = f[k];
    for (int c = 0; c < k; c++) s -= A[c][k] * y[c];
    y[k] = s / Rd[k];
  }
  y[6] = 0.0;
  // u = H0 H1 ... H5 y
  for (int k = 5; k >= 0; k--) {
    double s = 0.0;
    for (int r = k; r < 7; r++) s += A[r][k] * y[r];
    s *= beta[k];
    for (int r = k; r < 7; r++) y[r] -= s * A[r][k];
  }
  for (int r = 0; r < 7; r++) u[r] = y[r];
  return ok;
}

// mj_jacBody restricted to what MuJoCoHelper::GetJacobian keeps: rows [jacp; jacr], columns = dofs 0..6
inline void jac_body_6x7(const Model* m, const Data<double>* d, int body, double J[6][7]) {
  for (int r = 0; r < 6; r++)
    for (int c = 0; c < 7; c++) J[r][c] = 0.0;
  double offset[3];
  int root = m->body_rootid[body];
  for (int k = 0; k < 3; k++) offset[k] = d->xpos[3 * body + k] - d->subtree_com[3 * root + k];
  while (body > 0 && m->body_dofnum[body] == 0) body = m->body_parentid[body];
  if (body == 0) return;
  int i = m->body_dofadr[body] + m->body_dofnum[body] - 1;
  while (i >= 0) {
    if (i < 7) {
      const double* c = &d->cdof[6 * i];
      double tmp[3];
      cross(tmp, c, offset);
      for (int k = 0; k < 3; k++) { J[k][i] = c[3 + k] + tmp[k]; J[3 + k][i] = c[k]; }
    }
    i = m->dof_parentid[i];
  }
}

// CreateAllEETransitPoints (PushBaseClass.cpp:117-137): running sums, as written
inline std::vector<Pt3> all_transit_points(const std::vector<Pt3>& way, const std::vector<int>& timing) {
  std::vector<Pt3> path;
  path.push_back(way[0]);
  int counter = 1;
  for (size_t i = 0; i + 1 < way.size(); i++) {
    double dx = way[i + 1][0] - way[i][0], dy = way[i + 1][1] - way[i][1], dz = way[i + 1][2] - way[i][2];
    for (int j = 0; j < timing[i + 1]; j++) {
      Pt3 p;
      p[0] = path[counter - 1][0] + (dx / timing[i + 1]);
      p[1] = path[counter - 1][1] + (dy / timing[i + 1]);
      p[2] = path[counter - 1][2] + (dz / timing[i + 1]);
      path.push_back(p);
      counter++;
    }
  }
  return path;
}

// the second way-point of EEWayPointsSetup (PushBaseClass.cpp:8-44): 5 cm behind the object on the push line
inline Pt3 waypoint_setup(const double goal_xy[2], const double obj_pos[3]) {
  double angle = std::atan2(goal_xy[1] - obj_pos[1], goal_xy[0] - obj_pos[0]);
  Pt3 p;
  p[0] = obj_pos[0] - 0.05 * std::cos(angle);
  p[1] = obj_pos[1] - 0.05 * std::sin(angle);
  p[2] = 0.28f;
  return p;
}

// the second way-point of EEWayPointsPush (PushBaseClass.cpp:46-115)
inline Pt3 waypoint_push(const double goal_xy[2], const double obj_pos[3], int horizon, double timestep) {
  double angle = std::atan2(goal_xy[1] - obj_pos[1], goal_xy[0] - obj_pos[0]);
  double cylinder_radius = 0.01;
  double xo = cylinder_radius * std::cos(angle), yo = cylinder_radius * std::sin(angle);
  double ex = goal_xy[0] - xo, ey;
  if (goal_xy[1] - obj_pos[1] > 0) ey = goal_xy[1] + yo;
  else ey = goal_xy[1] - yo;
  double ix = obj_pos[0], iy = obj_pos[1];
  // 0.1 * ((5.0f/6.0f) * horizon * dt): float * int is evaluated in float
  double max_dist = 0.1 * ((5.0f / 6.0f) * horizon * timestep);
  double want = std::sqrt(std::pow((ex - ix), 2) + std::pow((ey - iy), 2));
  double prop = max_dist / want;
  Pt3 p;
  if (prop > 1) { p[0] = ex; p[1] = ey; }
  else { p[0] = ix + ((ex - ix) * prop); p[1] = iy + ((ey - iy) * prop); }
  p[2] = 0.28f;
  return p;
}

// JacobianEEControl (PushBaseClass.cpp:139-248) on `d` (= main_data).  U[len][m]; returns warning bits.
inline int jacobian_ee_control(const Model* m, const Task& task, Data<double>* d, const std::vector<Pt3>& path,
                               double EE_angle, double* U) {
  int warn = 0;
  const int ee = task.aux[1], nctrl = task.m;
  const SolverOpts opt = model_opts(m);
  EE_angle -= (kRefPI / 4);
  if (EE_angle < -(kRefPI / 2)) EE_angle = (2 * kRefPI) + EE_angle;
  double xa[3] = {std::cos(EE_angle), std::sin(EE_angle), 0}, za[3] = {0, 0, -1}, ya[3];
  ya[0] = (za[1] * xa[2]) - (za[2] * xa[1]);
  ya[1] = (za[2] * xa[0]) - (za[0] * xa[2]);
  ya[2] = (za[0] * xa[1]) - (za[1] * xa[0]);
  double rot[9] = {xa[0], ya[0], za[0], xa[1], ya[1], za[1], xa[2], ya[2], za[2]};
  double desired[4];
  rotmat2quat_ref(rot, desired);
  bool quaternion_check = false;
  const float gainsTorque[6] = {100, 100, 200, 80, 80, 80};
  // GetRobotJointsGravityCompensationControls: qfrc_bias from the dof of the robot's first joint on
  const int start_index = task.dof_adr[0];
  for (size_t pi = 0; pi < path.size(); pi++) {
    double cur[4] = {d->xquat[4 * ee], d->xquat[4 * ee + 1], d->xquat[4 * ee + 2], d->xquat[4 * ee + 3]};
    if (!quaternion_check) {
      quaternion_check = true;
      double dp = 0;
      for (int j = 0; j < 4; j++) dp += cur[j] * desired[j];
      if (dp < 0) for (int j = 0; j < 4; j++) desired[j] = -desired[j];
    }
    double inv[4] = {cur[0], -cur[1], -cur[2], -cur[3]}, qd[4], axis[3], diff[6], force[6];
    mult_quat_ref(desired, inv, qd);
    quat2axis_ref(qd, axis);
    for (int j = 0; j < 3; j++) { diff[j] = path[pi][j] - d->xpos[3 * ee + j]; diff[j + 3] = axis[j]; }
    double J[6][7], u[7];
    jac_body_6x7(m, d, ee, J);
    for (int j = 0; j < 6; j++) force[j] = diff[j] * gainsTorque[j];
    if (!min_norm_solve_6x7(J, force, u)) warn |= 64;
    forward_skip(m, d, STAGE_NONE, opt);  // mj_forward inside GetRobotJointsGravityCompensationControls
    for (int j = 0; j < nctrl; j++) {
      u[j] += d->qfrc_bias[start_index + j];
      U[pi * nctrl + j] = u[j];
      d->ctrl[task.ctrl_adr[j]] = u[j];  // SetControlVector
    }
    step_skip(m, d, STAGE_NONE, opt);
    warn |= d->warning;
  }
  return warn;
}

// GenTestingData.cpp:177-220 for one trial of BoxSweep (setup_horizon = 0: CreateInitSetupControls only runs
// mj_forward) or TwoDPushing (setup_horizon = 1000).  start = the snapshot InitialiseSystemToStartState leaves.
// Out: snapshot0 (what saved_systems_state_list[0] holds when Optimise starts), U[T][m], optionally U_setup.
inline int init_sequence_ee(const Model* m, const Task& task, const double* start, const double goal_xy[2],
                            int setup_horizon, int T, double* snapshot0, double* U, double* U_setup) {
  if (task.m != 7 || task.aux.size() < 2) throw std::runtime_error("init_sequence_ee: needs 7 controls and aux = {object body, EE body}");
  const int obj = task.aux[0], ee = task.aux[1];
  const int obj_qadr = m->jnt_qposadr[m->body_jntadr[obj]];
  const SolverOpts opt = model_opts(m);
  Data<double> master(m), main(m);
  load_snapshot(m, &master, start);
  cp_mjdata(&main, &master);
  step_skip(m, &master, STAGE_NONE, opt);  // :189
  int warn = master.warning;
  // CreateInitSetupControls
  cp_mjdata(&main, &master);
  forward_skip(m, &main, STAGE_NONE, opt);
  if (setup_horizon > 0) {
    std::vector<Pt3> way;
    std::vector<int> timing;
    way.push_back({main.xpos[3 * ee], main.xpos[3 * ee + 1], main.xpos[3 * ee + 2]});
    timing.push_back(0);
    way.push_back(waypoint_setup(goal_xy, &main.qpos[obj_qadr]));
    timing.push_back(setup_horizon - 1);
    std::vector<Pt3> path = all_transit_points(way, timing);
    double angle = std::atan2(goal_xy[1] - master.qpos[obj_qadr + 1], goal_xy[0] - master.qpos[obj_qadr]);
    std::vector<double> us((size_t)path.size() * task.m);
    warn |= jacobian_ee_control(m, task, &main, path, angle, us.data());
    if (U_setup) std::copy(us.begin(), us.end(), U_setup);
  }
  cp_mjdata(&master, &main);  // :200-205
  // CreateInitOptimisationControls
  {
    std::vector<Pt3> way;
    std::vector<int> timing;
    way.push_back({main.xpos[3 * ee], main.xpos[3 * ee + 1], main.xpos[3 * ee + 2]});
    timing.push_back(0);
    way.push_back(waypoint_push(goal_xy, &main.qpos[obj_qadr], T, m->timestep));
    timing.push_back(T - 1);
    std::vector<Pt3> path = all_transit_points(way, timing);
    double angle = std::atan2(goal_xy[1] - master.qpos[obj_qadr + 1], goal_xy[0] - master.qpos[obj_qadr]);
    warn |= jacobian_ee_control(m, task, &main, path, angle, U);
  }
  save_snapshot(m, &master, snapshot0);
  return warn;
}

}  // namespace ok
