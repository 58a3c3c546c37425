// Initial-control generators on the device.
//
// tokp_init_controls_gravcomp restates pandaReaching::CreateInitOptimisationControls
// (src/ModelTranslator/Reaching.cpp:269-296): for t = 0..T-1, mj_forward at the current state, control := qfrc_bias
// of the robot's joints (MuJoCoHelper::GetRobotJointsGravityCompensationControls, MuJoCoHelper.cpp:200-231),
// mj_step, record the control - a closed-loop rollout, serial in t, one world per trajectory of the batch.
// mj_forward leaves the state untouched but stores its solution in qacc_warmstart (mj_fwdConstraint), so the step that
// follows starts its solver from the answer at the same state.
#include <vector>

#include "step_dyn.cuh"
#include "tokp_internal.h"
#include "traj.h"

namespace tokp {

void pack_snapshot(const tokp_traj* tr, const double* s, double* d);
void unpack_snapshot(const tokp_traj* tr, const double* d, double* s);
int check_applied_zero(const tokp_traj* tr, const double* s);

template <int G>
__global__ void __launch_bounds__(128, 1) k_init_gravcomp(DM dm, TaskDev task, const double* __restrict__ snap0,
                                                           double* __restrict__ U, int* __restrict__ status, int B, int T,
                                                           int SD, int ecap, int wsd, int iterations, double tolerance) {
  extern __shared__ double smem[];
  tk::Grp<G, 0> g;
  constexpr int GPW = 32 / G;
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int b = (blockIdx.x * (blockDim.x >> 5) + warp) * GPW + lane / G;
  bool run = b < B;
  unsigned wm = __ballot_sync(0xffffffffu, run);
  if (!run) return;
  g.wmask = wm;
  g.bsync = 0;
  tk::WS w;
  tk::ws_init(w, smem + (size_t)(warp * GPW + lane / G) * wsd, dm, ecap, G);
  int nq = dm.nq, nv = dm.nv, nu = dm.nu, m = task.m;
  const double* s = snap0 + (size_t)b * SD;
  double time = s[0];
  for (int i = g.gl; i < nq; i += G) w.qpos[i] = s[1 + i];
  for (int i = g.gl; i < nv; i += G) {
    w.qvel[i] = s[1 + nq + i];
    w.qacc[i] = s[1 + nq + nv + i];
    w.warm[i] = s[1 + nq + 2 * nv + i];
  }
  for (int i = g.gl; i < nu; i += G) w.ctrl[i] = s[1 + nq + 3 * nv + i];
  g.sync();
  int st = 0;
  for (int t = 0; t < T; t++) {
    tk::step_smooth(dm, w, g, st);   // mj_forward: qfrc_bias ...
    g.sync();
    tk::fwd_constraint(dm, w, g, iterations, tolerance);  // ... and the solver, which also leaves qacc_warmstart = qacc
    g.sync();
    for (int i = g.gl; i < m; i += G) {
      double u = w.frc_bias[task.dof_adr[i]];
      w.ctrl[task.ctrl_adr[i]] = u;  // SetControlVector: unclamped in mjData, clamped inside the actuation stage
      U[((size_t)b * T + t) * m + i] = u;
    }
    g.sync();
    tk::step(dm, w, g, iterations, tolerance, time, st);
    g.sync();
  }
  if (g.gl == 0) status[b] = st;
}


// ---- Jacobian end-effector path follower -------------------------------------------------------------------------
// PushBaseClass::JacobianEEControl (src/ModelTranslator/PushBaseClass.cpp:139-248) driven the way
// GenTestingData::GenDataOpenloopOptimisation drives it for one trial (src/GenTestingData.cpp:177-220):
//   mj_step(master_reset) -> CreateInitSetupControls (BoxSweep.cpp:63-70: mj_forward only; TwoDPushing.cpp:225-258:
//   way-points to a spot 5 cm behind the object, `setup` controller steps) -> saved[0] <- state ->
//   CreateInitOptimisationControls (BoxSweep.cpp:72-99, TwoDPushing.cpp:260-289: push way-points, T controller steps).
// The reference reads the end-effector pose and the body Jacobian (mj_jacBody) from main_data BEFORE the mj_forward of
// GetRobotJointsGravityCompensationControls, i.e. from the kinematics of the previous mj_step: one step stale from
// the second path point on.  Here the pose is kept in registers from the previous position stage and the Jacobian is
// read from cdof / rootcom, which live in the persistent part of the workspace and still hold that stage's values.

constexpr double kRefPI = 3.14152;  // #define PI 3.14152 (include/StdInclude.h:13)

// quat2Axis (src/StdInclude/StdInclude.cpp:91-122)
__device__ inline void quat2axis_ref(const double q[4], double out[3]) {
  double w = q[0];
  if (w > 1) w = 1;
  double angle = 2 * acos(w);
  double s = sqrt(1 - (w * w));
  if (s < 0.001) { out[0] = q[1]; out[1] = q[2]; out[2] = q[3]; }
  else { out[0] = q[1] / s; out[1] = q[2] / s; out[2] = q[3] / s; }
  for (int k = 0; k < 3; k++) out[k] *= angle;
}

// Eigen::Quaterniond(Matrix3d) as used by rotMat2Quat (StdInclude.cpp:176-187); R row-major
__device__ inline void rotmat2quat_ref(const double R[9], double q[4]) {
  double t = R[0] + R[4] + R[8];
  if (t > 0.0) {
    t = sqrt(t + 1.0);
    q[0] = 0.5 * t;
    t = 0.5 / t;
    q[1] = (R[7] - R[5]) * t; q[2] = (R[2] - R[6]) * t; q[3] = (R[3] - R[1]) * t;
  } else {
    int i = 0;
    if (R[4] > R[0]) i = 1;
    if (R[8] > R[4 * i]) i = 2;
    int j = (i + 1) % 3, k = (j + 1) % 3;
    t = sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
    q[1 + i] = 0.5 * t;
    t = 0.5 / t;
    q[0] = (R[3 * k + j] - R[3 * j + k]) * t;
    q[1 + j] = (R[3 * j + i] + R[3 * i + j]) * t;
    q[1 + k] = (R[3 * k + i] + R[3 * i + k]) * t;
  }
}

// u = pinv(J) f for the 6 x 7 body Jacobian: Householder QR of J' (7 x 6), R' y = f, u = Q [y; 0] - the minimum-norm
// solution Eigen's completeOrthogonalDecomposition().pseudoInverse() gives for a full-rank J (PushBaseClass.cpp:218-227)
__device__ inline bool min_norm_solve_6x7(double A[7][6] /* J', destroyed */, const double f[6], double u[7]) {
  double beta[6], Rd[6];
  for (int k = 0; k < 6; k++) {
    double nrm = 0.0;
    for (int r = k; r < 7; r++) nrm += A[r][k] * A[r][k];
    nrm = sqrt(nrm);
    double alpha = A[k][k] > 0.0 ? -nrm : nrm;
    double v0 = A[k][k] - alpha;
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
  for (int k = 0; k < 6; k++) rmax = fmax(rmax, fabs(Rd[k]));
  bool ok = true;
  for (int k = 0; k < 6; k++)
    if (!(fabs(Rd[k]) > 6.0 * 2.220446049250313e-16 * rmax)) ok = false;
  double y[7];
  for (int k = 0; k < 6; k++) {
    double s = f[k];
    for (int c = 0; c < k; c++) s -= A[c][k] * y[c];
    y[k] = s / Rd[k];
  }
  y[6] = 0.0;
  for (int k = 5; k >= 0; k--) {
    double s = 0.0;
    for (int r = k; r < 7; r++) s += A[r][k] * y[r];
    s *= beta[k];
    for (int r = k; r < 7; r++) y[r] -= s * A[r][k];
  }
  for (int r = 0; r < 7; r++) u[r] = y[r];
  return ok;
}

struct EeArgs {
  const double* snap_start;  // [B][SD] InitialiseSystemToStartState snapshots (device layout)
  const double* goal_xy;     // [B][2] residual_list[0].target
  double* snap0_out;         // [B][SD] what saved_systems_state_list[0] holds when Optimise starts
  double* U;                 // [B][T][m]
  double* U_setup;           // [B][setup][m]
  int* status;               // [B]
  int B, T, SD, pre_steps, setup, ecap, wsd, iterations;
  double tolerance;
};

enum { ST_EE_RANK = 64 };  // rank-deficient end-effector Jacobian

template <int G>
__global__ void __launch_bounds__(128, 1) k_init_ee(DM dm, TaskDev task, EeArgs a) {
  extern __shared__ double smem[];
  tk::Grp<G, 0> g;
  constexpr int GPW = 32 / G;
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int b = (blockIdx.x * (blockDim.x >> 5) + warp) * GPW + lane / G;
  bool run = b < a.B;
  unsigned wm = __ballot_sync(0xffffffffu, run);
  if (!run) return;
  g.wmask = wm;
  g.bsync = 0;
  tk::WS w;
  tk::ws_init(w, smem + (size_t)(warp * GPW + lane / G) * a.wsd, dm, a.ecap, G);
  const int nq = dm.nq, nv = dm.nv, nu = dm.nu, m = task.m;
  const int obj = task.aux[0], ee = task.aux[1];
  const int obj_qadr = dm.jnt_qposadr[dm.body_jntadr[obj]];
  const int ee_root = dm.body_rootid[ee];
  int ee_dofbody = ee;
  while (ee_dofbody > 0 && dm.body_dofnum[ee_dofbody] == 0) ee_dofbody = dm.body_parentid[ee_dofbody];
  const int ee_lastdof = ee_dofbody > 0 ? dm.body_dofadr[ee_dofbody] + dm.body_dofnum[ee_dofbody] - 1 : -1;
  const int bias0 = task.dof_adr[0];  // GetRobotJointsGravityCompensationControls: from the robot's first joint on
  const double* s = a.snap_start + (size_t)b * a.SD;
  double time = s[0];
  for (int i = g.gl; i < nq; i += G) w.qpos[i] = s[1 + i];
  for (int i = g.gl; i < nv; i += G) {
    w.qvel[i] = s[1 + nq + i];
    w.qacc[i] = s[1 + nq + nv + i];
    w.warm[i] = s[1 + nq + 2 * nv + i];
  }
  for (int i = g.gl; i < nu; i += G) w.ctrl[i] = s[1 + nq + 3 * nv + i];
  g.sync();
  int st = 0;
  const double gx = a.goal_xy[2 * b], gy = a.goal_xy[2 * b + 1];
  // mj_step(master_reset_data) (GenTestingData.cpp:189)
  for (int k = 0; k < a.pre_steps; k++) { tk::step(dm, w, g, a.iterations, a.tolerance, time, st); g.sync(); }
  // end-effector pose of the last kinematics on main_data
  double ee_pos[3], ee_quat[4];
  auto forward = [&]() {  // mj_forward
    tk::step_position(dm, w, g, st);
    g.sync();
    for (int k = 0; k < 3; k++) ee_pos[k] = w.xpos[3 * ee + k];
    for (int k = 0; k < 4; k++) ee_quat[k] = w.xquat[4 * ee + k];
    g.sync();
    tk::step_velocity(dm, w, g);
    g.sync();
    tk::fwd_constraint(dm, w, g, a.iterations, a.tolerance);
    g.sync();
  };
  // CreateInitSetupControls: main <- master, mj_forward(main)
  forward();
  for (int phase = 0; phase < 2; phase++) {
    const int len = phase == 0 ? a.setup : a.T;
    if (phase == 1) {
      // master_reset <- main (GenTestingData.cpp:200-205): the state Optimise starts from
      double* o = a.snap0_out + (size_t)b * a.SD;
      if (g.gl == 0) o[0] = time;
      for (int i = g.gl; i < nq; i += G) o[1 + i] = w.qpos[i];
      for (int i = g.gl; i < nv; i += G) {
        o[1 + nq + i] = w.qvel[i];
        o[1 + nq + nv + i] = w.qacc[i];
        o[1 + nq + 2 * nv + i] = w.warm[i];
      }
      for (int i = g.gl; i < nu; i += G) o[1 + nq + 3 * nv + i] = w.ctrl[i];
    }
    if (len <= 0) continue;
    // way-points: EEWayPointsSetup (PushBaseClass.cpp:8-44) / EEWayPointsPush (:46-115)
    const double ox = w.qpos[obj_qadr], oy = w.qpos[obj_qadr + 1];
    const double push_angle = atan2(gy - oy, gx - ox);
    double wp[3];
    if (phase == 0) {
      wp[0] = ox - 0.05 * cos(push_angle);
      wp[1] = oy - 0.05 * sin(push_angle);
    } else {
      double cylinder_radius = 0.01;
      double xo = cylinder_radius * cos(push_angle), yo = cylinder_radius * sin(push_angle);
      double ex = gx - xo, ey = (gy - oy > 0) ? gy + yo : gy - yo;
      double max_dist = 0.1 * ((double)((5.0f / 6.0f) * (float)len) * dm.timestep);
      double want = sqrt((ex - ox) * (ex - ox) + (ey - oy) * (ey - oy));
      double prop = max_dist / want;
      if (prop > 1) { wp[0] = ex; wp[1] = ey; }
      else { wp[0] = ox + ((ex - ox) * prop); wp[1] = oy + ((ey - oy) * prop); }
    }
    wp[2] = (double)0.28f;
    // CreateAllEETransitPoints (:117-137): running sums of diff / timing
    double pt[3] = {ee_pos[0], ee_pos[1], ee_pos[2]}, inc[3];
    for (int k = 0; k < 3; k++) inc[k] = (wp[k] - pt[k]) / (double)(len - 1);
    // JacobianEEControl (:139-248)
    double EE_angle = push_angle - (kRefPI / 4);
    if (EE_angle < -(kRefPI / 2)) EE_angle = (2 * kRefPI) + EE_angle;
    double desired[4];
    {
      double xa[3] = {cos(EE_angle), sin(EE_angle), 0}, za[3] = {0, 0, -1}, ya[3];
      ya[0] = (za[1] * xa[2]) - (za[2] * xa[1]);
      ya[1] = (za[2] * xa[0]) - (za[0] * xa[2]);
      ya[2] = (za[0] * xa[1]) - (za[1] * xa[0]);
      double rot[9] = {xa[0], ya[0], za[0], xa[1], ya[1], za[1], xa[2], ya[2], za[2]};
      rotmat2quat_ref(rot, desired);
    }
    const double gains[6] = {100, 100, 200, 80, 80, 80};
    for (int pi = 0; pi < len; pi++) {
      if (pi > 0) for (int k = 0; k < 3; k++) pt[k] = pt[k] + inc[k];
      if (pi == 0) {
        double dp = 0;
        for (int j = 0; j < 4; j++) dp += ee_quat[j] * desired[j];
        if (dp < 0) for (int j = 0; j < 4; j++) desired[j] = -desired[j];
      }
      double qd[4], axis[3], force[6], u[7];
      {
        // multQuat(desired, invQuat(current)) (StdInclude.cpp:189-227)
        double wr = desired[0], xr = desired[1], yr = desired[2], zr = desired[3];
        double ws = ee_quat[0], xs = -ee_quat[1], ys = -ee_quat[2], zs = -ee_quat[3];
        qd[0] = (wr * ws) - (xr * xs) - (yr * ys) - (zr * zs);
        qd[1] = (ws * xr) + (wr * xs) + (yr * zs) - (zr * ys);
        qd[2] = (ws * yr) + (wr * ys) + (zr * xs) - (xr * zs);
        qd[3] = (ws * zr) + (wr * zs) + (xr * ys) - (yr * xs);
      }
      quat2axis_ref(qd, axis);
      for (int j = 0; j < 3; j++) { force[j] = (pt[j] - ee_pos[j]) * gains[j]; force[j + 3] = axis[j] * gains[j + 3]; }
      // mj_jacBody columns 0..6 from the (stale) cdof / subtree_com, transposed
      double At[7][6];
      for (int r = 0; r < 7; r++) for (int c = 0; c < 6; c++) At[r][c] = 0.0;
      double off[3];
      for (int k = 0; k < 3; k++) off[k] = ee_pos[k] - w.rootcom[3 * ee_root + k];
      for (int i = ee_lastdof; i >= 0; i = dm.dof_parentid[i]) {
        if (i >= 7) continue;
        const double* c = &w.cdof[6 * i];
        double tmp[3];
        tk::cross3(tmp, c, off);
        for (int k = 0; k < 3; k++) { At[i][k] = c[3 + k] + tmp[k]; At[i][3 + k] = c[k]; }
      }
      if (!min_norm_solve_6x7(At, force, u)) st |= ST_EE_RANK;
      forward();  // mj_forward inside GetRobotJointsGravityCompensationControls; refreshes ee_pos / ee_quat for the next point
      double* Uo = phase == 0 ? (a.U_setup ? a.U_setup + ((size_t)b * a.setup + pi) * m : nullptr) : a.U + ((size_t)b * a.T + pi) * m;
      for (int j = g.gl; j < m; j += G) {
        double v = u[j] + w.frc_bias[bias0 + j];
        if (Uo) Uo[j] = v;
        w.ctrl[task.ctrl_adr[j]] = v;  // SetControlVector
      }
      g.sync();
      tk::step(dm, w, g, a.iterations, a.tolerance, time, st);
      g.sync();
    }
  }
  int bits = 0;
  for (int k = 0; k < 7; k++) if (g.any((st >> k) & 1)) bits |= (1 << k);
  if (g.gl == 0) a.status[b] = bits;
}

}  // namespace tokp

using namespace tokp;

extern "C" int tokp_init_controls_gravcomp(tokp_traj* tr, const double* snapshot0, double* U_out, int* status_out) {
  if (!tr || !snapshot0 || !U_out) { set_error("tokp_init_controls_gravcomp: bad argument"); return TOKP_ERR_ARG; }
  const DM& dm = tr->dm;
  int B = tr->B, T = tr->T, m = tr->m;
  if (m > tr->n) { set_error("tokp_init_controls_gravcomp: more controls than state joints"); return TOKP_ERR_ARG; }
  std::vector<double> d((size_t)B * tr->SD);
  for (int b = 0; b < B; b++) {
    int rc = check_applied_zero(tr, snapshot0 + (size_t)b * tr->S);
    if (rc) return rc;
    pack_snapshot(tr, snapshot0 + (size_t)b * tr->S, d.data() + (size_t)b * tr->SD);
  }
  // one scratch allocation (snapshots | controls | status), released on every return path
  struct Scratch { void* p = nullptr; ~Scratch() { if (p) cudaFree(p); } } scratch;
  TOKP_CUDA(cudaMalloc(&scratch.p, d.size() * 8 + (size_t)B * T * m * 8 + (size_t)B * 4));
  double* d_snap = static_cast<double*>(scratch.p);
  double* d_U = d_snap + d.size();
  int* d_st = reinterpret_cast<int*>(d_U + (size_t)B * T * m);
  TOKP_CUDA(cudaMemcpyAsync(d_snap, d.data(), d.size() * 8, cudaMemcpyHostToDevice, tr->stream));
  // full constraint-row capacity: this is a one-off generator, not a hot loop
  int G = tr->G < 8 ? 8 : tr->G;
  LaunchPlan lp;
  int rc = plan_launch(dm, G, 1, &lp);
  if (rc) return rc;
  int warps = lp.warps > 4 ? 4 : lp.warps;
  int wpb = warps * lp.gpw;
  size_t smem = (size_t)lp.wsd * 8 * wpb;
  int blocks = (B + wpb - 1) / wpb;
#define LAUNCH_INIT(GG)                                                                                            \
  TOKP_CUDA(cudaFuncSetAttribute(k_init_gravcomp<GG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));    \
  k_init_gravcomp<GG><<<blocks, warps * 32, smem, tr->stream>>>(dm, tr->task, d_snap, d_U, d_st, B, T, tr->SD,     \
                                                                 lp.ecap, lp.wsd, dm.iterations, dm.tolerance);
  if (G == 8) { LAUNCH_INIT(8) } else if (G == 16) { LAUNCH_INIT(16) } else { LAUNCH_INIT(32) }
#undef LAUNCH_INIT
  g_launches++;
  TOKP_CUDA(cudaGetLastError());
  std::vector<int> hst(B);
  TOKP_CUDA(cudaMemcpyAsync(U_out, d_U, (size_t)B * T * m * 8, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaMemcpyAsync(hst.data(), d_st, (size_t)B * 4, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  if (status_out) for (int b = 0; b < B; b++) status_out[b] = hst[b];
  return TOKP_OK;
}

extern "C" int tokp_init_controls_ee(tokp_traj* tr, const double* snapshot_start, const double* goal_xy, int pre_steps,
                                     int setup_horizon, double* snapshot0_out, double* U_out, double* U_setup_out,
                                     int* status_out) {
  if (!tr || !snapshot_start || !goal_xy || !snapshot0_out || !U_out || pre_steps < 0 || setup_horizon < 0 || setup_horizon == 1 ||
      tr->T < 2) {
    set_error("tokp_init_controls_ee: bad argument");
    return TOKP_ERR_ARG;
  }
  const DM& dm = tr->dm;
  int B = tr->B, T = tr->T, m = tr->m;
  if (m != 7 || tr->n < 7 || dm.nv < 7 || tr->task.aux[0] <= 0 || tr->task.aux[1] <= 0) {
    set_error("tokp_init_controls_ee: needs a 7-joint arm on dofs 0..6 and aux = {object body, end-effector body}");
    return TOKP_ERR_ARG;
  }
  std::vector<double> d((size_t)B * tr->SD);
  for (int b = 0; b < B; b++) {
    int rc = check_applied_zero(tr, snapshot_start + (size_t)b * tr->S);
    if (rc) return rc;
    pack_snapshot(tr, snapshot_start + (size_t)b * tr->S, d.data() + (size_t)b * tr->SD);
  }
  // one scratch allocation: start snapshots | goals | snapshot0 | U | U_setup | status
  size_t n_snap = d.size(), n_goal = (size_t)2 * B, n_U = (size_t)B * T * m, n_Us = (size_t)B * setup_horizon * m;
  double* d_buf = nullptr;
  TOKP_CUDA(cudaMalloc(&d_buf, (2 * n_snap + n_goal + n_U + n_Us) * 8 + (size_t)B * 4));
  struct Guard { double* p; ~Guard() { cudaFree(p); } } guard{d_buf};
  EeArgs a;
  double* p = d_buf;
  a.snap_start = p; p += n_snap;
  a.goal_xy = p; p += n_goal;
  a.snap0_out = p; p += n_snap;
  a.U = p; p += n_U;
  a.U_setup = setup_horizon > 0 ? p : nullptr; p += n_Us;
  a.status = reinterpret_cast<int*>(p);
  a.B = B; a.T = T; a.SD = tr->SD; a.pre_steps = pre_steps; a.setup = setup_horizon;
  a.iterations = dm.iterations; a.tolerance = dm.tolerance;
  TOKP_CUDA(cudaMemcpyAsync((void*)a.snap_start, d.data(), n_snap * 8, cudaMemcpyHostToDevice, tr->stream));
  TOKP_CUDA(cudaMemcpyAsync((void*)a.goal_xy, goal_xy, n_goal * 8, cudaMemcpyHostToDevice, tr->stream));
  int G = tr->G < 8 ? 8 : tr->G;
  LaunchPlan lp;
  int rc = plan_launch(dm, G, 1, &lp);  // full constraint-row capacity: a one-off generator, not a hot loop
  if (rc) return rc;
  a.ecap = lp.ecap; a.wsd = lp.wsd;
  // spread the trials over the SMs: one warp per block while the batch is smaller than the device
  int warps = lp.warps > 4 ? 4 : lp.warps;
  while (warps > 1 && (B + warps * lp.gpw - 1) / (warps * lp.gpw) < 148) warps >>= 1;
  int wpb = warps * lp.gpw;
  size_t smem = (size_t)lp.wsd * 8 * wpb;
  int blocks = (B + wpb - 1) / wpb;
#define LAUNCH_EE(GG)                                                                                       \
  TOKP_CUDA(cudaFuncSetAttribute(k_init_ee<GG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));    \
  k_init_ee<GG><<<blocks, warps * 32, smem, tr->stream>>>(dm, tr->task, a);
  if (G == 8) { LAUNCH_EE(8) } else if (G == 16) { LAUNCH_EE(16) } else { LAUNCH_EE(32) }
#undef LAUNCH_EE
  g_launches++;
  TOKP_CUDA(cudaGetLastError());
  std::vector<int> hst(B);
  std::vector<double> hs0(n_snap);
  TOKP_CUDA(cudaMemcpyAsync(hs0.data(), a.snap0_out, n_snap * 8, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaMemcpyAsync(U_out, a.U, n_U * 8, cudaMemcpyDeviceToHost, tr->stream));
  if (U_setup_out && n_Us) TOKP_CUDA(cudaMemcpyAsync(U_setup_out, a.U_setup, n_Us * 8, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaMemcpyAsync(hst.data(), a.status, (size_t)B * 4, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  for (int b = 0; b < B; b++) unpack_snapshot(tr, hs0.data() + (size_t)b * tr->SD, snapshot0_out + (size_t)b * tr->S);
  if (status_out) for (int b = 0; b < B; b++) status_out[b] = hst[b];
  return TOKP_OK;
}
