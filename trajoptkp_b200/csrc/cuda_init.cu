// Initial-control generators on the device.
//
// tokp_init_controls_gravcomp restates pandaReaching::CreateInitOptimisationControls
// (src/ModelTranslator/Reaching.cpp:269-296): for t = 0..T-1, mj_forward at the current state, control := qfrc_bias
// of the robot's joints (MuJoCoHelper::GetRobotJointsGravityCompensationControls, MuJoCoHelper.cpp:200-231),
// mj_step, record the control - a closed-loop rollout, serial in t, one world per trajectory of the batch.
// mj_forward leaves the state and the warm start untouched, so "forward then step" is the smooth stage evaluated
// once for the bias and once more inside the step.
#include <vector>

#include "step_dyn.cuh"
#include "tokp_internal.h"
#include "traj.h"

namespace tokp {

void pack_snapshot(const tokp_traj* tr, const double* s, double* d);
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
    tk::step_smooth(dm, w, g, st);   // mj_forward up to qfrc_bias
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
  double *d_snap = nullptr, *d_U = nullptr;
  int* d_st = nullptr;
  TOKP_CUDA(cudaMalloc(&d_snap, d.size() * 8));
  TOKP_CUDA(cudaMalloc(&d_U, (size_t)B * T * m * 8));
  TOKP_CUDA(cudaMalloc(&d_st, (size_t)B * 4));
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
  cudaFree(d_snap); cudaFree(d_U); cudaFree(d_st);
  if (status_out) for (int b = 0; b < B; b++) status_out[b] = hst[b];
  return TOKP_OK;
}
