// Device arena of a trajectory batch + common CUDA helpers (internal).
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>
#include <string>
#include <vector>

#include "device_model.h"
#include "tokp_internal.h"

namespace tokp {

extern std::atomic<int64_t> g_launches;

#define TOKP_CUDA(expr)                                                                         \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      tokp::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));                      \
      return TOKP_ERR_CUDA;                                                                     \
    }                                                                                           \
  } while (0)

// returns TOKP_OK or TOKP_ERR_CUDA; uploads on first use
int ensure_device_model(const tokp_model* m, const DM** out);

// number of lanes per world for a model (8, 16 or 32)
int lanes_for_model(const DM& dm);
// TOKP_WARPS / TOKP_EFC_CAP into dm.opt_* (called where a DM copy is made, never on a launch path)
void read_env_options(DM& dm);

// Shared-memory launch plan of the world kernels.  pass 0 = fast pass with a small constraint-row capacity
// (TOKP_EFC_CAP, default min(maxefc, 32)); pass 1 = redo pass with ecap = maxefc for the worlds that overflowed.
struct LaunchPlan {
  int G, gpw, ecap, wsd, warps, worlds_per_block;
  int resident_blocks;  // blocks of this shape that stay resident on the whole device (persistent-grid size)
  size_t smem;
};
int fast_ecap(const DM& dm);
// kernel dispatch on (lanes per world, compile-time nv): the shipped models get a specialised instantiation
// (acrobot nv = 2, walker nv = 9), everything else the generic one (NV = 0 reads nv at run time)
#define TOKP_DISPATCH(G, NVV, LAUNCH)                                                   \
  if ((G) == 8 && (NVV) == 9) { LAUNCH(8, 9) }                                          \
  else if ((G) == 8 && (NVV) == 2) { LAUNCH(8, 2) }                                     \
  else if ((G) == 32 && (NVV) == 9) { LAUNCH(32, 9) }                                   \
  else if ((G) == 16 && (NVV) == 9) { LAUNCH(16, 9) }                                   \
  else if ((G) == 4) { LAUNCH(4, 0) }                                                   \
  else if ((G) == 8) { LAUNCH(8, 0) }                                                   \
  else if ((G) == 16) { LAUNCH(16, 0) }                                                 \
  else { LAUNCH(32, 0) }
// register budget of the world kernels: 128-thread blocks per SM that __launch_bounds__ asks for (2 -> 255 registers,
// 8 warps; 16-lane groups trade registers for residency: 3 -> 168 registers, 12 warps)
constexpr int reg_blocks(int lanes) { return lanes == 16 ? 3 : 2; }
inline bool needs_redo_pass(const DM& dm) { return fast_ecap(dm) < dm.maxefc; }
// returns TOKP_OK or TOKP_ERR_UNSUPPORTED (workspace does not fit shared memory)
int plan_launch(const DM& dm, int G, int pass, LaunchPlan* out);

// compact device snapshot: [time | qpos nq | qvel nv | qacc nv | warm nv | ctrl nu]
inline int dsnap_size(const DM& dm) { return 1 + dm.nq + 3 * dm.nv + dm.nu; }

struct TaskDev {
  int kind, n, m, nr;
  int ntg;      // target scalars per trajectory (nr unless the task has vector targets)
  int aux[4];   // model ids the residual reads (BoxSweep: goal body, gripper body)
  int kp_method, min_N, max_N;
  double iter_thr;
  const int *qpos_adr, *dof_adr, *ctrl_adr;           // device
  const double *weight, *weight_terminal, *jerk_thr, *vel_thr, *ctrl_limits;  // device
};

}  // namespace tokp

struct tokp_traj {
  const tokp_model* model = nullptr;
  DM dm;
  tokp::TaskDev task;
  int B = 0, T = 0, S = 0 /* host snapshot size */, SD = 0 /* device snapshot size */, W = 0 /* worlds per key-point */;
  int n = 0, m = 0, nx = 0, nr = 0, G = 8;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr, ev3 = nullptr;
  void* arena = nullptr;
  size_t arena_bytes = 0;
  // device arrays (all inside arena)
  double *snap, *X, *U, *resid, *A, *Bm, *rx, *ru, *lx, *lxx, *lu, *luu, *K, *k, *target;
  double *cand;       // [B][NA][T+1][SD] rollout buffers (zero-initialised, rows 0 and T never written)
  double *cand_cost;  // [B][NA]
  double *fd_out;     // [B][T][W][nq+nv]
  unsigned long long* kp_mask;  // [B][T]
  unsigned long long* kp_last;  // [B] velocity_change: scan hits of row T-1 (before every dof is appended again)
  int* kp_flags;      // [B]: bit0 = velocity_change duplicate last row
  int *kp_prev, *kp_next;  // [B][T][n]
  int* status;        // [B][T][W]
  int* wstatus;       // [B][2] sticky OR of the step status bits of every FD world / rollout world of trajectory b
  double *cost, *deltaJ, *lambda_d;  // [B]
  double *snap0_in, *U0_in;          // resident copies of the last host inputs of tokp_optimise ([B][SD], [B][T][m])
  int* bp_ok;         // [B]
  int* fd_list2;      // [B*T*W] FD worlds sorted for the lane kernel
  int *fd_list0, *fd_list1;  // [B*T*W] compacted FD world lists
  int* fd_counts;     // [1024] list sizes and cursors of the FD launch, bin sizes of the lane kernel (cuda_fd.cu, fd_lanes.cuh)
  double* fd_share = nullptr;  // [fd_share_rows][ShareLayout.total] stage-sharing slots (own allocation, made on first use)
  int fd_share_rows = 0, fd_share_slots_per_row = 1;
  int* fd_rowslot = nullptr;   // [fd_share_rows] slot of each row's nominal world (tail of the fd_share allocation)
  int fd_last_lane_mode = 0;
  // rollout tuning (TOKP_ROLL_BLOCKSYNC, TOKP_SPREAD_WARPS, TOKP_SPREAD_KEEP_LANES; read once at creation)
  int roll_blocksync = -1, spread_warps = 0, spread_keep_lanes = 0;
  int spread_blocksync = -1;  // Grp::bsync level of the spread (one world per warp) rollouts; -1 = the default (2)
  // key-point statistics without a bulk readback: k_kp_stats leaves [B][3] (sum of per-dof percentages, FD worlds,
  // columns) in kp_stats; it, the FD list counters and the status words land in pinned host memory with the stream
  double* kp_stats = nullptr;
  double* h_kp_stats = nullptr;  // pinned [B][3]
  int* h_fd_counts = nullptr;    // pinned [8]
  int* h_wstatus = nullptr;      // pinned [B][2]
  void* cols_scratch = nullptr;  // tokp_dynamics_derivs_cols: zero [B*T] mask + one row index (allocated on first use)
  void* ie_scratch = nullptr;    // iterative_error: need mask | row list | segments | verdicts (allocated on first use)
  int fd_share_gib = 16;       // cap of the lane_mode-2 share buffer (TOKP_FD_SHARE_GIB, read once at creation)
  int fd_stage_sharing = 1;    // mj_stepSkip-style stage reuse between the worlds of a key-point (tokp_traj_set_option)
  int fd_lanes = 2;            // lane-per-world FD kernel where the model allows it: 1 = the stage-reusing worlds, 2 = every
                               // world (the group kernel then runs the position stage only), 0 = off
  int sm_count = 148;
  int fd_blocksync = 2;        // Grp::bsync level of the FD kernels (TOKP_FD_BLOCKSYNC, read once at creation)
  int* roll_flags;    // [B][NA] rollout worlds that overflowed the fast constraint-row capacity
  int* task_ints;     // qpos_adr | dof_adr | ctrl_adr
  double* task_reals;
  static const int NA = 6;
  bool have_rollout = false, have_inputs = false, fd_timed = false;
  double fd_kernel_ms = 0, fd_kernel_ms_sum = 0;
  int64_t fd_kernel_launches = 0;
  double timers[3] = {0, 0, 0};
  int64_t counters[4] = {0, 0, 0, 0};
  std::vector<double> h_stage;  // host staging
  // Optimiser::filteringMethod / lowPassACoefficient / FIRCoefficients (include/Optimiser/Optimiser.h:173,219-220)
  int filter_method = 0;  // TOKP_FILTER_*
  double low_pass_a = 0.25;
  int nfir = 5;
  double fir[16] = {0.1, 0.15, 0.5, 0.15, 0.1};
};
