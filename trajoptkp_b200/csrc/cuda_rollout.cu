// Rollout kernels: nominal rollout and the alpha-batched closed-loop line-search rollouts, persistent over
// the whole horizon in one launch.  Replaces iLQR::RolloutTrajectory (src/Optimiser/iLQR.cpp:202-254) and
// iLQR::ForwardsPassParallel (:824-934) + Optimiser::SaveSystemStateToRolloutData (Optimiser.cpp:416-438).
#include <cstdlib>
#include "step_dyn.cuh"
#include "task_dev.cuh"
#include "traj.h"

namespace tokp {

struct RollArgs {
  double* snap;        // [B][T+1][SD] nominal snapshots
  double* X;           // [B][T+1][nx]
  double* U;           // [B][T][m]      U_old (written in nominal mode, read in forward mode)
  double* resid;       // [B][T+1][nr]   (written in nominal mode)
  const double* U_in;  // [B][T][m]      nominal mode: controls to apply
  const double* K;     // [B][T][m*nx]
  const double* k;     // [B][T][m]
  const double* target;  // [B][nr]
  double* cand;        // [B][NA][T+1][SD]
  double* cost;        // nominal: [B]; forward: [B][NA]
  const double* alphas;  // [NA]
  int B, T, SD, NA, mode;  // mode 0 nominal, 1 forward pass
  int iterations;
  double tolerance;
  TaskDev task;
  int* flags;          // [nworlds] overflow flags written by pass 0, consumed by pass 1
  int ecap, pass, last;  // last: no further pass re-runs overflowed worlds
  int spread;          // 1: one world per warp (first group only) - latency layout for small batches
  int bsync;           // Grp::bsync level: the block walks the stages of every step together
  int* wstatus;        // [B][2] sticky per-trajectory warning bits (column 1 = rollout worlds)
};

template <int G, int NV>
__global__ void __launch_bounds__(G == 16 ? 384 : 128, G == 16 ? 1 : 2) k_rollout(DM dm, RollArgs a, int wsd) {
  extern __shared__ double smem[];
  tk::Grp<G, NV> g;
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int GPW = 32 / G;
  int widx = a.spread ? blockIdx.x * (blockDim.x >> 5) + warp : (blockIdx.x * (blockDim.x >> 5) + warp) * GPW + lane / G;
  int slot = a.spread ? warp : warp * GPW + lane / G;
  int nworlds = a.mode == 0 ? a.B : a.B * a.NA;
  // block-synchronous launches keep every thread: groups past the end repeat the last world (same values to the
  // same addresses)
  if (a.bsync && widx >= nworlds) widx = nworlds - 1;
  bool run = widx < nworlds && !(a.spread && lane / G != 0);
  if (run && a.pass == 1 && !a.flags[widx]) run = false;
  unsigned wm = __ballot_sync(0xffffffffu, run);
  if (!run) return;
  g.wmask = wm;
  g.bsync = a.bsync;
  int b = a.mode == 0 ? widx : widx / a.NA;
  int ai = a.mode == 0 ? 0 : widx - b * a.NA;
  tk::WS w;
  tk::ws_init(w, smem + (size_t)slot * wsd, dm, a.ecap, G);
  const TaskDev& tk_ = a.task;
  int nq = dm.nq, nv = dm.nv, nu = dm.nu, n = tk_.n, m = tk_.m, nx = 2 * n, nr = tk_.nr, T = a.T, SD = a.SD;
  double* snapb = a.snap + (size_t)b * (T + 1) * SD;
  const double* s = snapb;
  double time = s[0];
  for (int i = g.gl; i < nq; i += G) w.qpos[i] = s[1 + i];
  for (int i = g.gl; i < nv; i += G) {
    w.qvel[i] = s[1 + nq + i];
    w.qacc[i] = s[1 + nq + nv + i];
    w.warm[i] = s[1 + nq + 2 * nv + i];
  }
  for (int i = g.gl; i < nu; i += G) w.ctrl[i] = s[1 + nq + 3 * nv + i];
  g.sync();
  double* Xb = a.X + (size_t)b * (T + 1) * nx;
  double* Ub = a.U + (size_t)b * T * m;
  const double* tgt = a.target + (size_t)b * tk_.ntg;
  double cost = 0.0;
  int st = 0;
  double alpha = a.mode ? a.alphas[ai] : 0.0;
  double* candb = a.cand + ((size_t)b * a.NA + ai) * (T + 1) * SD;
  if (a.mode == 0) {
    for (int i = g.gl; i < n; i += G) { Xb[i] = w.qpos[tk_.qpos_adr[i]]; Xb[n + i] = w.qvel[tk_.dof_adr[i]]; }
  }
  for (int t = 0; t < T; t++) {
    if (a.mode == 0) {
      for (int i = g.gl; i < m; i += G) w.ctrl[tk_.ctrl_adr[i]] = a.U_in[((size_t)b * T + t) * m + i];
      g.sync();
    } else {
      // U_new = U_old[t] + alpha k[t] + K[t] (x - X_old[t]), clamped (iLQR.cpp:847-890)
      const double* Kt = a.K + ((size_t)b * T + t) * m * nx;
      const double* kt = a.k + ((size_t)b * T + t) * m;
      const double* Xo = Xb + (size_t)t * nx;
      for (int i = g.gl; i < m; i += G) {
        double gsum = 0.0;
        for (int c = 0; c < nx; c++) {
          double xc = c < n ? w.qpos[tk_.qpos_adr[c]] : w.qvel[tk_.dof_adr[c - n]];
          gsum += Kt[c * m + i] * (xc - Xo[c]);
        }
        double u = Ub[(size_t)t * m + i] + (alpha * kt[i]) + gsum;
        if (u > tk_.ctrl_limits[2 * i + 1]) u = tk_.ctrl_limits[2 * i + 1];
        if (u < tk_.ctrl_limits[2 * i]) u = tk_.ctrl_limits[2 * i];
        w.tmpv[i] = u;
      }
      g.sync();
      for (int i = g.gl; i < m; i += G) w.ctrl[tk_.ctrl_adr[i]] = w.tmpv[i];
      g.sync();
      // cost of (x_t, u_t) before the step (iLQR.cpp:896-912)
      if (g.gl == 0) {
        double r[kMaxResid];
        task_residuals(dm, tk_, w.qpos, w.qvel, w.ctrl, tgt, r);
        cost = __dadd_rn(cost, task_cost(tk_, r, t == T - 1));
      }
    }
    tk::step(dm, w, g, a.iterations, a.tolerance, time, st);
    if (a.mode == 0) {
      // residuals / cost after the step, X_old, U_old, saved[t+1] (iLQR.cpp:222-246)
      if (g.gl == 0) {
        double r[kMaxResid];
        task_residuals(dm, tk_, w.qpos, w.qvel, w.ctrl, tgt, r);
        double* rb = a.resid + ((size_t)b * (T + 1) + t) * nr;
        for (int i = 0; i < nr; i++) rb[i] = r[i];
        cost = __dadd_rn(cost, task_cost(tk_, r, t == T - 1));
      }
      double* Xn = Xb + (size_t)(t + 1) * nx;
      for (int i = g.gl; i < n; i += G) { Xn[i] = w.qpos[tk_.qpos_adr[i]]; Xn[n + i] = w.qvel[tk_.dof_adr[i]]; }
      for (int i = g.gl; i < m; i += G) Ub[(size_t)t * m + i] = w.ctrl[tk_.ctrl_adr[i]];
    }
    if (a.mode == 0 || t != 0) {
      double* o = a.mode == 0 ? snapb + (size_t)(t + 1) * SD : candb + (size_t)t * SD;
      if (g.gl == 0) o[0] = time;
      for (int i = g.gl; i < nq; i += G) o[1 + i] = w.qpos[i];
      for (int i = g.gl; i < nv; i += G) {
        o[1 + nq + i] = w.qvel[i];
        o[1 + nq + nv + i] = w.qacc[i];
        o[1 + nq + 2 * nv + i] = w.warm[i];
      }
      for (int i = g.gl; i < nu; i += G) o[1 + nq + 3 * nv + i] = w.ctrl[i];
    }
    g.sync();
  }
  int ovf = g.any(st & tk::ST_OVERFLOW);
  if (a.last && ovf) cost = __longlong_as_double(0x7ff8000000000000ll);  // constraint rows beyond the capacity: loud
  if (g.gl == 0) a.cost[a.mode == 0 ? b : b * a.NA + ai] = cost;
  if (g.gl == 0) a.flags[widx] = (a.pass == 0 && ovf) ? 1 : 0;
  int bits = 0;
  for (int k = 0; k < 6; k++) if (g.any((st >> k) & 1)) bits |= (1 << k);
  if (!a.last) bits &= ~tk::ST_OVERFLOW;
  if (g.gl == 0 && bits) atomicOr(&a.wstatus[2 * b + 1], bits);
}

int launch_rollout(tokp_traj* tr, int mode, const double* d_Uin, const double* d_alphas, int nalpha) {
  const DM& dm = tr->dm;
  RollArgs a;
  a.snap = tr->snap; a.X = tr->X; a.U = tr->U; a.resid = tr->resid; a.U_in = d_Uin; a.K = tr->K; a.k = tr->k;
  a.target = tr->target; a.cand = tr->cand; a.cost = mode == 0 ? tr->cost : tr->cand_cost; a.alphas = d_alphas;
  a.B = tr->B; a.T = tr->T; a.SD = tr->SD; a.NA = mode == 0 ? 1 : nalpha; a.mode = mode;
  a.iterations = dm.iterations; a.tolerance = dm.tolerance;
  a.task = tr->task;
  if (mode == 1) a.NA = tokp_traj::NA;  // buffer stride; unused alpha slots are skipped by the grid size
  int G = tr->G;
  a.flags = tr->roll_flags; a.wstatus = tr->wstatus;
  int nworlds = mode == 0 ? tr->B : tr->B * tokp_traj::NA;
  // few worlds: latency layout - one world per warp with all 32 lanes on it (fewer trips through every strided
  // loop), one warp per block so that the worlds spread over the SMs
  a.spread = nworlds <= 148 * 4;
  if (a.spread && !tr->spread_keep_lanes) G = 32;
  for (int pass = 0; pass < (needs_redo_pass(dm) ? 2 : 1); pass++) {
    LaunchPlan lp;
    int rc = plan_launch(dm, G, pass, &lp);
    if (rc) return rc;
    a.ecap = lp.ecap; a.pass = pass; a.last = (pass == 1 || !needs_redo_pass(dm)) ? 1 : 0;
    int warps = lp.warps;
    int blocks = (nworlds + lp.worlds_per_block - 1) / lp.worlds_per_block;
    if (a.spread) {
      // one world per warp; as many warps per block as it takes to cover the worlds with one block per SM (at most 4),
      // so that the warps of an SM walk the step together
      warps = (nworlds + 147) / 148;
      if (warps > 4) warps = 4;
      if (tr->spread_warps) warps = tr->spread_warps;
      while (warps > 1 && (size_t)lp.wsd * 8 * warps > 226 * 1024) warps--;  // large models: fewer worlds per block
      blocks = (nworlds + warps - 1) / warps;
      lp.smem = (size_t)lp.wsd * 8 * warps;
    }
    // the redo pass runs a sparse subset of the worlds: no block walk; neither for one-warp blocks
    a.bsync = (pass == 0 && (!a.spread || (warps > 1 && G == 32))) ? 2 : 0;
    if (tr->roll_blocksync >= 0 && a.bsync) a.bsync = tr->roll_blocksync;
    if (a.spread && tr->spread_blocksync >= 0 && a.bsync) a.bsync = tr->spread_blocksync;
#define LAUNCH(GG, NN)                                                                                              \
  TOKP_CUDA(cudaFuncSetAttribute(k_rollout<GG, NN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lp.smem));    \
  k_rollout<GG, NN><<<blocks, warps * 32, lp.smem, tr->stream>>>(dm, a, lp.wsd);
    TOKP_DISPATCH(G, dm.spec_nv, LAUNCH)
#undef LAUNCH
    g_launches++;
  }
  TOKP_CUDA(cudaGetLastError());
  return TOKP_OK;
}

}  // namespace tokp
