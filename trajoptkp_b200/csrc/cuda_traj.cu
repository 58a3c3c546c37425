// Trajectory batch: device arena, snapshot I/O, named array readback.
// Mirrors the storage the reference keeps in Optimiser (include/Optimiser/Optimiser.h:194-211),
// iLQR::Resize (src/Optimiser/iLQR.cpp:37-200) and MuJoCoHelper::saved_systems_state_list
// (include/MuJoCoHelper.h:118-123), as flat [B][T(+1)][...] device arrays.
#include <cstdlib>
#include <cstring>

#include "traj.h"

using namespace tokp;

namespace {

struct ArenaPlan {
  size_t bytes = 0;
  size_t take(size_t n, size_t elem) {
    size_t off = bytes;
    bytes += ((n * elem + 255) / 256) * 256;
    return off;
  }
};

}  // namespace

extern "C" int tokp_traj_create(const tokp_model* m, const tokp_task_desc* td, int horizon, int batch, tokp_traj** out) {
  if (!m || !td || !out || horizon < 2 || batch < 1) { set_error("tokp_traj_create: bad argument"); return TOKP_ERR_ARG; }
  if (td->n < 1 || td->n > 64 || td->m < 0 || td->num_resid < 1) { set_error("tokp_traj_create: bad task sizes"); return TOKP_ERR_ARG; }
  const DM* dmp;
  int rc = ensure_device_model(m, &dmp);
  if (rc) return rc;
  auto* tr = new tokp_traj();
  tr->model = m;
  tr->dm = *dmp;
  const DM& dm = tr->dm;
  int B = batch, T = horizon, n = td->n, mm = td->m, nx = 2 * n, nr = td->num_resid;
  tr->B = B; tr->T = T; tr->n = n; tr->m = mm; tr->nx = nx; tr->nr = nr;
  tr->S = tokp_snapshot_size(m);
  tr->SD = dsnap_size(dm);
  tr->W = 1 + 2 * mm + 4 * n;
  tr->G = lanes_for_model(dm);
  for (int i = 0; i < n; i++)
    if (td->qpos_adr[i] < 0 || td->qpos_adr[i] >= dm.nq || td->dof_adr[i] < 0 || td->dof_adr[i] >= dm.nv) {
      set_error("tokp_traj_create: state address out of range"); delete tr; return TOKP_ERR_ARG;
    }
  for (int i = 0; i < mm; i++)
    if (td->ctrl_adr[i] < 0 || td->ctrl_adr[i] >= dm.nu) { set_error("tokp_traj_create: ctrl address out of range"); delete tr; return TOKP_ERR_ARG; }
  ArenaPlan p;
  size_t NA = tokp_traj::NA;
  size_t o_snap = p.take((size_t)B * (T + 1) * tr->SD, 8);
  size_t o_X = p.take((size_t)B * (T + 1) * nx, 8);
  size_t o_U = p.take((size_t)B * T * mm, 8);
  size_t o_resid = p.take((size_t)B * (T + 1) * nr, 8);
  size_t o_A = p.take((size_t)B * T * nx * nx, 8);
  size_t o_B = p.take((size_t)B * T * nx * mm, 8);
  size_t o_rx = p.take((size_t)B * (T + 1) * nr * nx, 8);
  size_t o_ru = p.take((size_t)B * (T + 1) * nr * mm, 8);
  size_t o_lx = p.take((size_t)B * (T + 1) * nx, 8);
  size_t o_lxx = p.take((size_t)B * (T + 1) * nx * nx, 8);
  size_t o_lu = p.take((size_t)B * (T + 1) * mm, 8);
  size_t o_luu = p.take((size_t)B * (T + 1) * mm * mm, 8);
  size_t o_K = p.take((size_t)B * T * mm * nx, 8);
  size_t o_k = p.take((size_t)B * T * mm, 8);
  const int ntg = td->num_target > 0 ? td->num_target : nr;
  size_t o_target = p.take((size_t)B * ntg, 8);
  size_t o_cand = p.take((size_t)B * NA * (T + 1) * tr->SD, 8);
  size_t o_candcost = p.take((size_t)B * NA, 8);
  size_t o_fd = p.take((size_t)B * T * tr->W * (dm.nq + dm.nv), 8);
  size_t o_mask = p.take((size_t)B * T, 8);
  size_t o_last = p.take((size_t)B, 8);
  size_t o_flags = p.take((size_t)B, 4);
  size_t o_prev = p.take((size_t)B * T * n, 4);
  size_t o_next = p.take((size_t)B * T * n, 4);
  size_t o_status = p.take((size_t)B * T * tr->W, 4);
  size_t o_wst = p.take((size_t)B * 2, 4);
  size_t o_cost = p.take((size_t)B, 8);
  size_t o_dJ = p.take((size_t)B, 8);
  size_t o_lam = p.take((size_t)B, 8);
  size_t o_ok = p.take((size_t)B, 4);
  size_t o_rflags = p.take((size_t)B * NA, 4);
  size_t o_l0 = p.take((size_t)B * T * tr->W, 4);
  size_t o_l1 = p.take((size_t)B * T * tr->W, 4);
  size_t o_l2 = p.take((size_t)B * T * tr->W, 4);
  size_t o_cnt = p.take(1024, 4);
  size_t o_s0 = p.take((size_t)B * tr->SD, 8);
  size_t o_u0 = p.take((size_t)B * T * mm, 8);
  size_t o_tints = p.take((size_t)(2 * n + mm), 4);
  size_t o_treals = p.take((size_t)(2 * nr + 2 * n + 2 * mm), 8);
  cudaError_t e = cudaMalloc(&tr->arena, p.bytes);
  if (e != cudaSuccess) { set_error(std::string("traj arena: ") + cudaGetErrorString(e)); delete tr; return TOKP_ERR_CUDA; }
  tr->arena_bytes = p.bytes;
  cudaMemset(tr->arena, 0, p.bytes);
  uint8_t* base = (uint8_t*)tr->arena;
  tr->snap = (double*)(base + o_snap); tr->X = (double*)(base + o_X); tr->U = (double*)(base + o_U);
  tr->resid = (double*)(base + o_resid); tr->A = (double*)(base + o_A); tr->Bm = (double*)(base + o_B);
  tr->rx = (double*)(base + o_rx); tr->ru = (double*)(base + o_ru);
  tr->lx = (double*)(base + o_lx); tr->lxx = (double*)(base + o_lxx); tr->lu = (double*)(base + o_lu); tr->luu = (double*)(base + o_luu);
  tr->K = (double*)(base + o_K); tr->k = (double*)(base + o_k); tr->target = (double*)(base + o_target);
  tr->cand = (double*)(base + o_cand); tr->cand_cost = (double*)(base + o_candcost); tr->fd_out = (double*)(base + o_fd);
  tr->kp_mask = (unsigned long long*)(base + o_mask); tr->kp_last = (unsigned long long*)(base + o_last); tr->kp_flags = (int*)(base + o_flags);
  tr->kp_prev = (int*)(base + o_prev); tr->kp_next = (int*)(base + o_next); tr->status = (int*)(base + o_status);
  tr->wstatus = (int*)(base + o_wst);
  tr->cost = (double*)(base + o_cost); tr->deltaJ = (double*)(base + o_dJ); tr->lambda_d = (double*)(base + o_lam);
  tr->bp_ok = (int*)(base + o_ok);
  tr->roll_flags = (int*)(base + o_rflags);
  tr->fd_list0 = (int*)(base + o_l0); tr->fd_list1 = (int*)(base + o_l1); tr->fd_counts = (int*)(base + o_cnt);
  tr->fd_list2 = (int*)(base + o_l2);
  tr->snap0_in = (double*)(base + o_s0); tr->U0_in = (double*)(base + o_u0);
  tr->task_ints = (int*)(base + o_tints); tr->task_reals = (double*)(base + o_treals);
  // task tables
  std::vector<int> ti;
  ti.insert(ti.end(), td->qpos_adr, td->qpos_adr + n);
  ti.insert(ti.end(), td->dof_adr, td->dof_adr + n);
  ti.insert(ti.end(), td->ctrl_adr, td->ctrl_adr + mm);
  std::vector<double> trl;
  trl.insert(trl.end(), td->resid_weight, td->resid_weight + nr);
  trl.insert(trl.end(), td->resid_weight_terminal, td->resid_weight_terminal + nr);
  trl.insert(trl.end(), td->jerk_thresholds, td->jerk_thresholds + n);
  trl.insert(trl.end(), td->velchange_thresholds, td->velchange_thresholds + n);
  // ReturnControlLimits: first 2m entries of actuator_ctrlrange (MuJoCoHelper.cpp:262-264)
  for (int i = 0; i < 2 * mm; i++) trl.push_back(m->host.actuator_ctrlrange[i]);
  cudaMemcpy(tr->task_ints, ti.data(), ti.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(tr->task_reals, trl.data(), trl.size() * 8, cudaMemcpyHostToDevice);
  TaskDev& t = tr->task;
  t.kind = td->kind; t.n = n; t.m = mm; t.nr = nr; t.ntg = ntg;
  for (int i = 0; i < 4; i++) t.aux[i] = (td->aux && i < td->num_aux) ? td->aux[i] : -1;
  t.kp_method = td->kp_method; t.min_N = td->min_N; t.max_N = td->max_N; t.iter_thr = td->iterative_error_threshold;
  t.qpos_adr = tr->task_ints; t.dof_adr = tr->task_ints + n; t.ctrl_adr = tr->task_ints + 2 * n;
  t.weight = tr->task_reals; t.weight_terminal = tr->task_reals + nr; t.jerk_thr = tr->task_reals + 2 * nr;
  t.vel_thr = tr->task_reals + 2 * nr + n; t.ctrl_limits = tr->task_reals + 2 * nr + 2 * n;
  // every trajectory starts with the YAML targets
  std::vector<double> tg((size_t)B * ntg);
  for (int b = 0; b < B; b++) std::memcpy(&tg[(size_t)b * ntg], td->resid_target, ntg * 8);
  cudaMemcpy(tr->target, tg.data(), tg.size() * 8, cudaMemcpyHostToDevice);
  cudaMalloc(&tr->kp_stats, (size_t)B * 3 * 8);
  cudaMallocHost(&tr->h_kp_stats, (size_t)B * 3 * 8);
  cudaMallocHost(&tr->h_fd_counts, 8 * sizeof(int));
  cudaMallocHost(&tr->h_wstatus, (size_t)B * 2 * sizeof(int));
  if (!tr->kp_stats || !tr->h_kp_stats || !tr->h_fd_counts || !tr->h_wstatus) { set_error("tokp_traj_create: statistics buffers"); tokp_traj_destroy(tr); return TOKP_ERR_CUDA; }
  read_env_options(tr->dm);
  if (const char* e = getenv("TOKP_ROLL_BLOCKSYNC")) tr->roll_blocksync = atoi(e) > 3 ? 3 : (atoi(e) < 0 ? 0 : atoi(e));
  if (const char* e = getenv("TOKP_SPREAD_BLOCKSYNC")) tr->spread_blocksync = atoi(e) > 3 ? 3 : (atoi(e) < 0 ? 0 : atoi(e));
  if (const char* e = getenv("TOKP_SPREAD_WARPS")) { int v = atoi(e); if (v >= 1 && v <= 4) tr->spread_warps = v; }
  if (getenv("TOKP_SPREAD_KEEP_LANES")) tr->spread_keep_lanes = 1;
  if (const char* e = getenv("TOKP_FD_SHARE_GIB")) { int v = atoi(e); if (v >= 1 && v <= 64) tr->fd_share_gib = v; }
  if (const char* e = getenv("TOKP_FD_BLOCKSYNC")) tr->fd_blocksync = atoi(e) < 0 ? 0 : (atoi(e) > 3 ? 3 : atoi(e));
  {
    int dev = 0, smc = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&smc, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && smc > 0) tr->sm_count = smc;
  }
  cudaStreamCreate(&tr->stream);
  cudaEventCreate(&tr->ev0);
  cudaEventCreate(&tr->ev1);
  cudaEventCreate(&tr->ev2);
  cudaEventCreate(&tr->ev3);
  e = cudaGetLastError();
  if (e != cudaSuccess) { set_error(std::string("traj init: ") + cudaGetErrorString(e)); tokp_traj_destroy(tr); return TOKP_ERR_CUDA; }
  *out = tr;
  return TOKP_OK;
}

extern "C" void tokp_traj_destroy(tokp_traj* tr) {
  if (!tr) return;
  if (tr->stream) { cudaStreamSynchronize(tr->stream); cudaStreamDestroy(tr->stream); }
  if (tr->ev0) cudaEventDestroy(tr->ev0);
  if (tr->ev1) cudaEventDestroy(tr->ev1);
  if (tr->ev2) cudaEventDestroy(tr->ev2);
  if (tr->ev3) cudaEventDestroy(tr->ev3);
  if (tr->arena) cudaFree(tr->arena);
  if (tr->fd_share) cudaFree(tr->fd_share);
  if (tr->kp_stats) cudaFree(tr->kp_stats);
  if (tr->h_kp_stats) cudaFreeHost(tr->h_kp_stats);
  if (tr->h_fd_counts) cudaFreeHost(tr->h_fd_counts);
  if (tr->h_wstatus) cudaFreeHost(tr->h_wstatus);
  if (tr->cols_scratch) cudaFree(tr->cols_scratch);
  if (tr->ie_scratch) cudaFree(tr->ie_scratch);
  delete tr;
}

extern "C" int tokp_traj_set_targets(tokp_traj* tr, int b, const double* resid_target) {
  if (!tr || !resid_target || b < 0 || b >= tr->B) { set_error("tokp_traj_set_targets: bad argument"); return TOKP_ERR_ARG; }
  TOKP_CUDA(cudaMemcpyAsync(tr->target + (size_t)b * tr->task.ntg, resid_target, tr->task.ntg * 8, cudaMemcpyHostToDevice, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  return TOKP_OK;
}

extern "C" int tokp_traj_set_keypoint_method(tokp_traj* tr, int kp_method, int min_N, int max_N, const double* jerk_thr,
                                             const double* vel_thr, double iter_thr) {
  if (!tr || kp_method < 0 || kp_method > 4 || min_N < 1 || max_N < 1) { set_error("tokp_traj_set_keypoint_method: bad argument"); return TOKP_ERR_ARG; }
  tr->task.kp_method = kp_method; tr->task.min_N = min_N; tr->task.max_N = max_N; tr->task.iter_thr = iter_thr;
  if (jerk_thr) TOKP_CUDA(cudaMemcpy((void*)tr->task.jerk_thr, jerk_thr, tr->n * 8, cudaMemcpyHostToDevice));
  if (vel_thr) TOKP_CUDA(cudaMemcpy((void*)tr->task.vel_thr, vel_thr, tr->n * 8, cudaMemcpyHostToDevice));
  return TOKP_OK;
}

// tuning / test switches of a trajectory batch
extern "C" int tokp_traj_set_option(tokp_traj* tr, const char* name, int value) {
  if (!tr || !name) { set_error("tokp_traj_set_option: bad argument"); return TOKP_ERR_ARG; }
  std::string nm(name);
  if (nm == "fd_stage_sharing") { tr->fd_stage_sharing = value ? 1 : 0; return TOKP_OK; }
  if (nm == "fd_lanes") { tr->fd_lanes = value < 0 ? 0 : (value > 2 ? 2 : value); return TOKP_OK; }
  if (nm == "fd_blocksync") { tr->fd_blocksync = value < 0 ? 0 : (value > 3 ? 3 : value); return TOKP_OK; }
  set_error("tokp_traj_set_option: unknown option " + nm);
  return TOKP_ERR_ARG;
}

// Per-trajectory step warnings (the mjWARN_* analogues): out[b][0] = OR over every FD world of trajectory b since the
// last clear, out[b][1] = the same for rollout worlds (nominal + line search).  Bits: 1 bad qpos, 2 bad qvel, 4 bad
// qacc, 8 constraint rows beyond TOKP_EFC_MAX (contacts dropped), 32 unsupported collision pair / condim.
extern "C" int tokp_traj_status(tokp_traj* tr, int* out, int clear) {
  if (!tr || !out) { set_error("tokp_traj_status: bad argument"); return TOKP_ERR_ARG; }
  TOKP_CUDA(cudaMemcpyAsync(out, tr->wstatus, (size_t)tr->B * 2 * 4, cudaMemcpyDeviceToHost, tr->stream));
  if (clear) TOKP_CUDA(cudaMemsetAsync(tr->wstatus, 0, (size_t)tr->B * 2 * 4, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  return TOKP_OK;
}

namespace tokp {
// after a phase: unsupported features and dropped contacts make A / B or a rollout wrong - fail loudly instead of
// handing the optimiser a wrong model (the reference's MuJoCo would have aborted in mj_loadXML / raised mjWARN_CNSTRFULL)
int check_world_status(tokp_traj* tr, int column, const char* what) {
  std::vector<int> h((size_t)tr->B * 2);
  TOKP_CUDA(cudaMemcpyAsync(h.data(), tr->wstatus, h.size() * 4, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  for (int b = 0; b < tr->B; b++) {
    int bits = h[2 * b + column];
    if (bits & 32) { set_error(std::string(what) + ": a world hit an unsupported collision pair or contact dimension"); return TOKP_ERR_UNSUPPORTED; }
    if (column == 0 && (bits & 8)) {
      set_error(std::string(what) + ": constraint rows of an FD world exceed TOKP_EFC_MAX, contacts were dropped (trajectory " +
                std::to_string(b) + ")");
      return TOKP_ERR_STATE;
    }
  }
  return TOKP_OK;
}

// host snapshot [S] <-> device snapshot [SD]
void pack_snapshot(const tokp_traj* tr, const double* s, double* d) {
  const DM& dm = tr->dm;
  int nq = dm.nq, nv = dm.nv;
  d[0] = s[0];
  std::memcpy(d + 1, s + 1, (nq + 3 * nv) * 8);
  std::memcpy(d + 1 + nq + 3 * nv, s + tr->S - dm.nu, dm.nu * 8);
}
void unpack_snapshot(const tokp_traj* tr, const double* d, double* s) {
  const DM& dm = tr->dm;
  int nq = dm.nq, nv = dm.nv;
  std::memset(s, 0, tr->S * 8);
  s[0] = d[0];
  std::memcpy(s + 1, d + 1, (nq + 3 * nv) * 8);
  std::memcpy(s + tr->S - dm.nu, d + 1 + nq + 3 * nv, dm.nu * 8);
}
int check_applied_zero(const tokp_traj* tr, const double* s) {
  const DM& dm = tr->dm;
  const double* p = s + 1 + dm.nq + 3 * dm.nv;
  for (int i = 0; i < dm.nv + 6 * dm.nbody; i++)
    if (p[i] != 0.0) { set_error("non-zero qfrc_applied/xfrc_applied is outside the hot-path scope"); return TOKP_ERR_UNSUPPORTED; }
  return TOKP_OK;
}
}  // namespace tokp

extern "C" int tokp_traj_set_snapshot(tokp_traj* tr, int b, int t, const double* snapshot) {
  if (!tr || !snapshot || b < 0 || b >= tr->B || t < 0 || t > tr->T) { set_error("tokp_traj_set_snapshot: bad argument"); return TOKP_ERR_ARG; }
  int rc = check_applied_zero(tr, snapshot);
  if (rc) return rc;
  std::vector<double> d(tr->SD);
  pack_snapshot(tr, snapshot, d.data());
  TOKP_CUDA(cudaMemcpyAsync(tr->snap + ((size_t)b * (tr->T + 1) + t) * tr->SD, d.data(), tr->SD * 8, cudaMemcpyHostToDevice, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  return TOKP_OK;
}

extern "C" int tokp_traj_get_snapshot(tokp_traj* tr, int b, int t, double* snapshot) {
  if (!tr || !snapshot || b < 0 || b >= tr->B || t < 0 || t > tr->T) { set_error("tokp_traj_get_snapshot: bad argument"); return TOKP_ERR_ARG; }
  std::vector<double> d(tr->SD);
  TOKP_CUDA(cudaMemcpyAsync(d.data(), tr->snap + ((size_t)b * (tr->T + 1) + t) * tr->SD, tr->SD * 8, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  unpack_snapshot(tr, d.data(), snapshot);
  return TOKP_OK;
}

namespace {
// (pointer, per-trajectory element count) of a named device array
bool named_array(tokp_traj* tr, const std::string& nm, double** p, size_t* cnt) {
  int T = tr->T, nx = tr->nx, m = tr->m, nr = tr->nr;
#define F(name, ptr, c) if (nm == name) { *p = tr->ptr; *cnt = (size_t)(c); return true; }
  F("X_old", X, (T + 1) * nx) F("U_old", U, T * m) F("resid", resid, (T + 1) * nr)
  F("A", A, (size_t)T * nx * nx) F("B", Bm, (size_t)T * nx * m)
  F("r_x", rx, (size_t)(T + 1) * nr * nx) F("r_u", ru, (size_t)(T + 1) * nr * m)
  F("l_x", lx, (T + 1) * nx) F("l_xx", lxx, (size_t)(T + 1) * nx * nx) F("l_u", lu, (T + 1) * m) F("l_uu", luu, (T + 1) * m * m)
  F("K", K, (size_t)T * m * nx) F("k", k, T * m) F("target", target, tr->task.ntg)
  F("snap", snap, (size_t)(T + 1) * tr->SD) F("cand", cand, (size_t)tokp_traj::NA * (T + 1) * tr->SD)
  F("fd_out", fd_out, (size_t)T * tr->W * (tr->dm.nq + tr->dm.nv))
#undef F
  return false;
}
}  // namespace

extern "C" int tokp_traj_get(tokp_traj* tr, int b, const char* name, double* out, int cap) {
  if (!tr || !name || b < 0 || b >= tr->B) return -1000000;
  double* p;
  size_t cnt;
  if (!named_array(tr, name, &p, &cnt)) return -1000000;
  if ((size_t)cap < cnt || !out) return -(int)cnt;
  if (cudaMemcpyAsync(out, p + (size_t)b * cnt, cnt * 8, cudaMemcpyDeviceToHost, tr->stream) != cudaSuccess) return -1000000;
  cudaStreamSynchronize(tr->stream);
  return (int)cnt;
}

extern "C" int tokp_traj_set(tokp_traj* tr, int b, const char* name, const double* in, int count) {
  if (!tr || !name || !in || b < 0 || b >= tr->B) { set_error("tokp_traj_set: bad argument"); return TOKP_ERR_ARG; }
  double* p;
  size_t cnt;
  if (!named_array(tr, name, &p, &cnt) || (size_t)count != cnt) { set_error("tokp_traj_set: unknown array or wrong size"); return TOKP_ERR_ARG; }
  TOKP_CUDA(cudaMemcpyAsync(p + (size_t)b * cnt, in, cnt * 8, cudaMemcpyHostToDevice, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  return TOKP_OK;
}

extern "C" int tokp_traj_counters(tokp_traj* tr, int64_t out[4]) {
  if (!tr || !out) return TOKP_ERR_ARG;
  for (int i = 0; i < 4; i++) out[i] = tr->counters[i];
  return TOKP_OK;
}

extern "C" int tokp_fd_kernel_ms(tokp_traj* tr, double out[3]) {
  if (!tr || !out) return TOKP_ERR_ARG;
  if (tr->fd_timed) {
    cudaEventSynchronize(tr->ev3);
    float ms = 0;
    cudaEventElapsedTime(&ms, tr->ev2, tr->ev3);
    tr->fd_kernel_ms = ms; tr->fd_kernel_ms_sum += ms; tr->fd_kernel_launches++;
    tr->fd_timed = false;
  }
  out[0] = tr->fd_kernel_ms; out[1] = tr->fd_kernel_ms_sum; out[2] = (double)tr->fd_kernel_launches;
  return TOKP_OK;
}

extern "C" int tokp_timers(tokp_traj* tr, double ms[3]) {
  if (!tr || !ms) return TOKP_ERR_ARG;
  for (int i = 0; i < 3; i++) ms[i] = tr->timers[i];
  return TOKP_OK;
}
