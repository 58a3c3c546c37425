// Device model upload + batched plain stepping (tokp_step_worlds).
// Replaces mj_makeData/mj_step usage of MuJoCoHelper (src/PhysicsSimulators/MuJoCoHelper.cpp:877-923)
// for batches of independent worlds.
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

#include "step_dyn.cuh"
#include "traj.h"

namespace tokp {

std::atomic<int64_t> g_launches{0};
static std::mutex g_upload_mutex;

void device_model_free(DeviceModel* dm) {
  if (!dm) return;
  if (dm->arena) cudaFree(dm->arena);
  delete dm;
}

void read_env_options(DM& dm) {
  dm.opt_warps = 0; dm.opt_efc_cap = 0;
  if (const char* e = getenv("TOKP_WARPS")) { int v = atoi(e); if (v > 0) dm.opt_warps = v; }
  if (const char* e = getenv("TOKP_EFC_CAP")) { int v = atoi(e); if (v > 0) dm.opt_efc_cap = v; }
}

// called at model upload / tokp_traj_create only (reads TOKP_LANES)
int lanes_for_model(const DM& dm) {
  if (const char* e = getenv("TOKP_LANES")) {  // tuning override (8 / 16 / 32)
    int v = atoi(e);
    if (v == 4 || v == 8 || v == 16 || v == 32) return v;
  }
  // 16 lanes from nv = 8: half the worlds per warp of the 8-lane layout, but 168 registers instead of 255 and half
  // the shared memory per warp, i.e. twice the resident warps (walker: 12 per SM)
  int k = dm.nv > dm.nbody ? dm.nv : dm.nbody;
  if (k < 8) return 8;
  if (k <= 20) return 16;
  return 32;
}

int fast_ecap(const DM& dm) {
  int cap = dm.fast_ecap > 0 ? dm.fast_ecap : 32;
  if (dm.opt_efc_cap > 0) cap = dm.opt_efc_cap;
  return cap < dm.maxefc ? cap : dm.maxefc;
}

int plan_launch(const DM& dm, int G, int pass, LaunchPlan* out) {
  LaunchPlan p;
  p.G = G; p.gpw = 32 / G;
  p.ecap = pass == 0 ? fast_ecap(dm) : dm.maxefc;
  p.wsd = tk::ws_doubles(dm, p.ecap, G);
  // the groups of a warp access their workspaces at the same offsets: a stride of 64 B (mod 128 B) between consecutive
  // workspaces spreads a unit-stride fp64 access of the whole warp over all 32 banks (2 wavefronts, the minimum)
  if (p.gpw > 1) while (p.wsd % 16 != 8) p.wsd++;
  size_t per_world = (size_t)p.wsd * 8;
  const int reg_warps = 4 * reg_blocks(G);  // warps per SM the register budget allows
  const size_t sm_bytes = 227 * 1024, block_overhead = 1024;
  // warps per block: two blocks per SM that together hold as many warps as shared memory and registers allow.  The
  // kernels walk the step block-synchronously (Grp::bsync), so the warps of a block share the instruction cache; two
  // such blocks per SM measured faster than one (barrier waits overlap) and faster than three or four.
  int best = 0;
  {
    size_t per_warp = per_world * p.gpw;
    int total = (int)((sm_bytes - 2 * block_overhead) / per_warp);
    if (total > reg_warps) total = reg_warps;
    best = total / 2;
    if (best < 1 && per_warp <= sm_bytes) best = 1;
  }
  if (dm.opt_warps > 0 && dm.opt_warps <= (G == 16 ? 12 : 4) && per_world * dm.opt_warps * p.gpw <= sm_bytes) best = dm.opt_warps;
  if (best == 0) { set_error("model too large for the shared-memory workspace"); return TOKP_ERR_UNSUPPORTED; }
  p.warps = best; p.worlds_per_block = best * p.gpw; p.smem = per_world * best * p.gpw;
  {
    static int sm_counts[64] = {0};  // per device
    int dev = 0;
    cudaGetDevice(&dev);
    int& sm_count = sm_counts[dev & 63];
    if (!sm_count &&
        (cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sm_count <= 0))
      sm_count = 148;
    int resident = (int)(sm_bytes / (p.smem + block_overhead));
    if (resident > 32) resident = 32;
    if (resident * best > reg_warps) resident = reg_warps / best;
    if (resident < 1) resident = 1;
    p.resident_blocks = resident * sm_count;
  }
  *out = p;
  return TOKP_OK;
}

int ensure_device_model(const tokp_model* mc, const DM** out) {
  tokp_model* m = const_cast<tokp_model*>(mc);
  std::lock_guard<std::mutex> lock(g_upload_mutex);
  if (m->dev) {
    int cur = -1;
    cudaGetDevice(&cur);
    if (cur != m->dev->device) {  // the cached pointers belong to another GPU
      set_error("tokp_model was uploaded to CUDA device " + std::to_string(m->dev->device) + ", the current device is " +
                std::to_string(cur) + ": load the model once per device");
      return TOKP_ERR_STATE;
    }
    *out = &m->dev->dm;
    return TOKP_OK;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    set_error("no CUDA device available (tokp has no CPU fallback)");
    return TOKP_ERR_CUDA;
  }
  const Model& h = m->host;
  // pack every array into one host buffer (8-byte aligned), upload, then patch pointers
  std::vector<uint8_t> buf;
  auto add = [&](const void* p, size_t bytes) {
    size_t off = buf.size();
    buf.resize(off + ((bytes + 7) / 8) * 8 + 8);
    if (bytes) std::memcpy(buf.data() + off, p, bytes);
    return off;
  };
  struct Patch { size_t off; const void** dst; };
  std::vector<Patch> patches;
  auto* d = new DeviceModel();
  DM& dm = d->dm;
#define AI(name) patches.push_back({add(h.name.data(), h.name.size() * 4), (const void**)&dm.name});
#define AR(name) patches.push_back({add(h.name.data(), h.name.size() * 8), (const void**)&dm.name});
  AI(body_parentid) AI(body_rootid) AI(body_weldid) AI(body_jntadr) AI(body_jntnum) AI(body_dofadr) AI(body_dofnum)
  AI(body_depth) AI(body_subtreemask_lo) AI(body_subtreemask_hi)
  AI(jnt_type) AI(jnt_qposadr) AI(jnt_dofadr) AI(jnt_bodyid) AI(jnt_limited)
  AI(dof_bodyid) AI(dof_jntid) AI(dof_parentid) AI(dof_ancmask_lo) AI(dof_ancmask_hi) AI(dof_treelast)
  AI(geom_type) AI(geom_bodyid) AI(geom_condim) AI(geom_priority) AI(pair_geom1) AI(pair_geom2) AI(pair_slotadr) AI(geom_meshadr) AI(geom_meshnum)
  AI(actuator_trnid) AI(actuator_dofadr) AI(actuator_ctrllimited) AI(actuator_forcelimited) AI(actuator_biastype)
  AR(qpos0) AR(body_pos) AR(body_quat) AR(body_ipos) AR(body_iquat) AR(body_mass) AR(body_inertia) AR(body_invweight0)
  AR(body_gravcomp)
  AR(jnt_pos) AR(jnt_axis) AR(jnt_range) AR(jnt_solref) AR(jnt_solimp) AR(jnt_margin)
  AR(dof_armature) AR(dof_damping) AR(dof_invweight0)
  AR(geom_size) AR(geom_pos) AR(geom_quat) AR(geom_friction) AR(geom_solref) AR(geom_solimp) AR(geom_margin) AR(geom_gap)
  AR(geom_solmix) AR(geom_rbound) AR(mesh_vert)
  AR(actuator_gear) AR(actuator_ctrlrange) AR(actuator_forcerange) AR(actuator_gainprm) AR(actuator_biasprm)
#undef AI
#undef AR
  cudaError_t e = cudaMalloc(&d->arena, buf.size());
  if (e == cudaSuccess) e = cudaMemcpy(d->arena, buf.data(), buf.size(), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    set_error(std::string("model upload: ") + cudaGetErrorString(e));
    if (d->arena) cudaFree(d->arena);
    delete d;
    return TOKP_ERR_CUDA;
  }
  for (auto& p : patches) *p.dst = (const uint8_t*)d->arena + p.off;
  dm.nq = h.nq(); dm.nv = h.nv(); dm.nu = h.nu(); dm.nbody = h.nbody(); dm.njnt = h.njnt(); dm.ngeom = h.ngeom();
  dm.npair = h.npair(); dm.maxcontact = h.maxcontact() > 0 ? h.maxcontact() : 1; dm.maxefc = h.maxefc() > 0 ? h.maxefc() : 1;
  // The compiler's bound assumes every filtered pair in contact at once (Franka scenes: 345-949 rows); the redo pass
  // sizes its shared-memory workspace by this number, so it is capped like MuJoCo's njmax (TOKP_EFC_MAX, default
  // 192 rows).  A world that still overflows keeps ST_OVERFLOW in its status / gets a NaN rollout cost.
  {
    int cap = 192;
    if (const char* e = getenv("TOKP_EFC_MAX")) { int v = atoi(e); if (v > 0) cap = v; }
    if (dm.maxefc > cap) dm.maxefc = cap;
  }
  dm.treedepth = h.dims[10];
  dm.fast_ecap = 0;
  read_env_options(dm);
  d->lanes = 0;
  dm.iterations = h.opt_int[0]; dm.ls_iterations = h.opt_int[1];
  dm.timestep = h.opt_real[0];
  dm.gravity[0] = h.opt_real[1]; dm.gravity[1] = h.opt_real[2]; dm.gravity[2] = h.opt_real[3];
  dm.tolerance = h.opt_real[4]; dm.ls_tolerance = h.opt_real[5]; dm.meaninertia = h.opt_real[7];
  // the nv-specialised instantiations (TOKP_DISPATCH) compile the cylinder / mesh / convex colliders out: only models
  // whose every pair is plane-{sphere, capsule, box} may use them, whatever their nv
  dm.spec_nv = dm.nv;
  for (size_t p = 0; p < h.pair_geom1.size(); p++) {
    int t1 = h.geom_type[h.pair_geom1[p]], t2 = h.geom_type[h.pair_geom2[p]];
    if (!(t1 == GEOM_PLANE && (t2 == GEOM_SPHERE || t2 == GEOM_CAPSULE || t2 == GEOM_BOX))) dm.spec_nv = 0;
  }
  dm.max_tree_dofs = 0;
  for (int d = 0; d < dm.nv; d++) {  // a tree's first dof has no parent
    if (h.dof_parentid[d] < 0 && h.dof_treelast[d] - d + 1 > dm.max_tree_dofs) dm.max_tree_dofs = h.dof_treelast[d] - d + 1;
  }
  dm.any_damping = 0; dm.any_gravcomp = 0;
  for (double v : h.dof_damping) if (v > 0) dm.any_damping = 1;
  for (double v : h.body_gravcomp) if (v != 0) dm.any_gravcomp = 1;
  if (const char* e = getenv("TOKP_SPEC_NV")) { if (atoi(e) == 0) dm.spec_nv = 0; }  // tuning: generic kernels only
  dm.lane_ok = (dm.spec_nv == 2 || dm.spec_nv == 9) && !dm.any_gravcomp && dm.maxefc <= 64;
  for (int jt : h.jnt_type) if (jt != JNT_HINGE && jt != JNT_SLIDE) dm.lane_ok = 0;
  cudaGetDevice(&d->device);
  m->dev = d;
  *out = &m->dev->dm;
  return TOKP_OK;
}

// ---- tokp_step_worlds ---------------------------------------------------------------------------
// in/out: host-layout snapshots [count][S] already on the device
template <int G, int NV>
__global__ void k_step_worlds(DM dm, const double* __restrict__ in, double* __restrict__ out, int* __restrict__ status,
                              int count, int nsteps, int iterations, double tolerance, int S, int wsd, int ecap, int pass,
                              int forward_after) {
  extern __shared__ double smem[];
  constexpr int GPW = 32 / G;  // groups per warp
  tk::Grp<G, NV> g;
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int gidx = (blockIdx.x * (blockDim.x >> 5) + warp) * GPW + lane / G;  // world index
  int slot = warp * GPW + lane / G;
  bool run = gidx < count;
  if (run && pass == 1 && !(status[gidx] & tk::ST_OVERFLOW)) run = false;  // redo pass: only overflowed worlds
  unsigned wm = __ballot_sync(0xffffffffu, run);
  if (!run) return;
  g.wmask = wm;
  tk::WS w;
  tk::ws_init(w, smem + (size_t)slot * wsd, dm, ecap, G);
  const double* s = in + (size_t)gidx * S;
  int nq = dm.nq, nv = dm.nv;
  double time = s[0];
  for (int i = g.gl; i < nq; i += G) w.qpos[i] = s[1 + i];
  for (int i = g.gl; i < nv; i += G) {
    w.qvel[i] = s[1 + nq + i];
    w.qacc[i] = s[1 + nq + nv + i];
    w.warm[i] = s[1 + nq + 2 * nv + i];
  }
  for (int i = g.gl; i < dm.nu; i += G) w.ctrl[i] = s[S - dm.nu + i];
  g.sync();
  int st = 0;
#ifdef TOKP_STAGE_CLOCKS
  w.clk_on = gidx == 0;
  for (int k = 0; k < nsteps; k++) {
    w.clk = clock64();
    TK_CNT(w, g, 22);
    tk::step(dm, w, g, iterations, tolerance, time, st);
  }
#else
  for (int k = 0; k < nsteps; k++) tk::step(dm, w, g, iterations, tolerance, time, st);
#endif
  if (forward_after) {  // mj_forward: the state stays, qacc and qacc_warmstart become the solution at this state
    tk::step_smooth(dm, w, g, st);
    g.sync();
    tk::fwd_constraint(dm, w, g, iterations, tolerance);
    g.sync();
  }
  double* o = out + (size_t)gidx * S;
  for (int i = g.gl; i < S; i += G) o[i] = 0.0;
  g.sync();
  if (g.gl == 0) o[0] = time;
  for (int i = g.gl; i < nq; i += G) o[1 + i] = w.qpos[i];
  for (int i = g.gl; i < nv; i += G) {
    o[1 + nq + i] = w.qvel[i];
    o[1 + nq + nv + i] = w.qacc[i];
    o[1 + nq + 2 * nv + i] = w.warm[i];
  }
  for (int i = g.gl; i < dm.nu; i += G) o[S - dm.nu + i] = w.ctrl[i];
  // OR-reduce the status bits across the group
  int bits = 0;
  for (int b = 0; b < 6; b++) if (g.any((st >> b) & 1)) bits |= (1 << b);
  if (g.gl == 0) status[gidx] = bits;
}

}  // namespace tokp

using namespace tokp;

namespace tokp {
static int step_worlds_impl(const tokp_model* m, const double* snapshots_in, double* snapshots_out, int count, int nsteps,
                            int solver_iterations, double solver_tolerance, int* status, int forward_after) {
  if (!m || !snapshots_in || !snapshots_out || count <= 0 || nsteps < 0) { set_error("tokp_step_worlds: bad argument"); return TOKP_ERR_ARG; }
  const DM* dmp;
  int rc = ensure_device_model(m, &dmp);
  if (rc) return rc;
  const DM& dm = *dmp;
  int S = tokp_snapshot_size(m);
  // applied forces must be zero (outside the hot-path scope)
  for (int wi = 0; wi < count; wi++) {
    const double* s = snapshots_in + (size_t)wi * S + 1 + dm.nq + 3 * dm.nv;
    for (int i = 0; i < dm.nv + 6 * dm.nbody; i++)
      if (s[i] != 0.0) { set_error("tokp_step_worlds: non-zero qfrc_applied/xfrc_applied unsupported"); return TOKP_ERR_UNSUPPORTED; }
  }
  size_t bytes = (size_t)count * S * sizeof(double);
  // one scratch allocation (in | out | status), released on every return path
  struct Scratch { void* p = nullptr; ~Scratch() { if (p) cudaFree(p); } } scratch;
  TOKP_CUDA(cudaMalloc(&scratch.p, 2 * bytes + (size_t)count * sizeof(int)));
  double* d_in = static_cast<double*>(scratch.p);
  double* d_out = d_in + (size_t)count * S;
  int* d_st = reinterpret_cast<int*>(d_out + (size_t)count * S);
  TOKP_CUDA(cudaMemcpy(d_in, snapshots_in, bytes, cudaMemcpyHostToDevice));
  int iters = solver_iterations < 0 ? dm.iterations : solver_iterations;
  double tol = solver_iterations < 0 ? dm.tolerance : solver_tolerance;
  if (!m->dev->lanes) m->dev->lanes = lanes_for_model(dm);
  int G = m->dev->lanes;
  for (int pass = 0; pass < (needs_redo_pass(dm) ? 2 : 1); pass++) {
    LaunchPlan lp;
    rc = plan_launch(dm, G, pass, &lp);
    if (rc) return rc;
    int blocks = (count + lp.worlds_per_block - 1) / lp.worlds_per_block;
#define LAUNCH(GG, NN)                                                                                                    \
  TOKP_CUDA(cudaFuncSetAttribute(k_step_worlds<GG, NN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lp.smem));      \
  k_step_worlds<GG, NN><<<blocks, lp.warps * 32, lp.smem>>>(dm, d_in, d_out, d_st, count, nsteps, iters, tol, S, lp.wsd, \
                                                        lp.ecap, pass, forward_after);
    TOKP_DISPATCH(G, dm.spec_nv, LAUNCH)
#undef LAUNCH
    g_launches++;
  }
  TOKP_CUDA(cudaGetLastError());
  TOKP_CUDA(cudaDeviceSynchronize());
  TOKP_CUDA(cudaMemcpy(snapshots_out, d_out, bytes, cudaMemcpyDeviceToHost));
  if (status) TOKP_CUDA(cudaMemcpy(status, d_st, count * sizeof(int), cudaMemcpyDeviceToHost));
  return TOKP_OK;
}
}  // namespace tokp

extern "C" int tokp_step_worlds(const tokp_model* m, const double* snapshots_in, double* snapshots_out, int count,
                                int nsteps, int solver_iterations, double solver_tolerance, int* status) {
  return step_worlds_impl(m, snapshots_in, snapshots_out, count, nsteps, solver_iterations, solver_tolerance, status, 0);
}

#ifdef TOKP_STAGE_CLOCKS
// development builds only: cycles per stage of world 0 of the tokp_step_worlds launches so far (tools/stage_clocks.py)
extern "C" int tokp_stage_clocks(unsigned long long* out, int clear) {
  TOKP_CUDA(cudaMemcpyFromSymbol(out, tk::g_stage_clk, 32 * sizeof(unsigned long long)));
  if (clear) { unsigned long long z[32] = {0}; TOKP_CUDA(cudaMemcpyToSymbol(tk::g_stage_clk, z, sizeof(z))); }
  return TOKP_OK;
}
#endif

extern "C" int tokp_forward_worlds(const tokp_model* m, const double* snapshots_in, double* snapshots_out, int count,
                                   int nsteps, int solver_iterations, double solver_tolerance, int* status) {
  return step_worlds_impl(m, snapshots_in, snapshots_out, count, nsteps, solver_iterations, solver_tolerance, status, 1);
}

// ---- tokp_step_worlds_static: the same batched stepping without any allocation -------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch tokp_step_worlds takes
// from cudaMalloc (input snapshots | output snapshots | status words) lives here in a file-scope device array of
// TOKP_STATIC_SCRATCH_BYTES = 4 MiB, which every GPU that loads the library holds until the process exits.  All calls
// share the array and it keeps the last call's contents: calls that overlap on the GPU race (serialise them).
#define TOKP_STATIC_SCRATCH_BYTES (4u << 20)
static __device__ __align__(256) unsigned char g_static_scratch[TOKP_STATIC_SCRATCH_BYTES];

namespace tokp {
// In host code `g_static_scratch` is a host-side stand-in, not a device address: the address a kernel may use comes
// from cudaGetSymbolAddress, looked up once per device with that device current.
static int static_scratch_ptr(unsigned char** out) {
  constexpr int kMaxDev = 64;
  static unsigned char* ptrs[kMaxDev] = {nullptr};
  static std::mutex mu;
  int dev = -1;
  TOKP_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDev) { set_error("tokp_step_worlds_static: device index out of range"); return TOKP_ERR_CUDA; }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) {
      set_error(std::string("tokp_step_worlds_static: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
      return TOKP_ERR_CUDA;
    }
    ptrs[dev] = static_cast<unsigned char*>(p);
  }
  *out = ptrs[dev];
  return TOKP_OK;
}
}  // namespace tokp

extern "C" int tokp_step_worlds_static(const tokp_model* m, const double* snapshots_in, double* snapshots_out, int count,
                                       int nsteps, int solver_iterations, double solver_tolerance, int* status) {
  if (!m || !snapshots_in || !snapshots_out || count <= 0 || nsteps < 0) { set_error("tokp_step_worlds_static: bad argument"); return TOKP_ERR_ARG; }
  // the model upload allocates: it must have happened already (tokp_traj_create or tokp_step_worlds on this model)
  if (!m->dev) {
    set_error("tokp_step_worlds_static: the model is not on the device yet (call tokp_traj_create or tokp_step_worlds once first; "
              "this entry point allocates nothing)");
    return TOKP_ERR_STATE;
  }
  const DM& dm = m->dev->dm;
  const int S = tokp_snapshot_size(m);
  const size_t snap_bytes = (size_t)count * S * sizeof(double);
  const size_t need = 2 * snap_bytes + (size_t)count * sizeof(int);
  if (need > TOKP_STATIC_SCRATCH_BYTES) {
    set_error("tokp_step_worlds_static: " + std::to_string(count) + " worlds of " + std::to_string(S) + " doubles need " +
              std::to_string(need) + " bytes of scratch, the static scratch holds " + std::to_string(TOKP_STATIC_SCRATCH_BYTES) +
              " (at most " + std::to_string(TOKP_STATIC_SCRATCH_BYTES / (2 * (size_t)S * sizeof(double) + sizeof(int))) +
              " worlds of this model per call; use tokp_step_worlds for larger batches)");
    return TOKP_ERR_ARG;
  }
  for (int wi = 0; wi < count; wi++) {
    const double* s = snapshots_in + (size_t)wi * S + 1 + dm.nq + 3 * dm.nv;
    for (int i = 0; i < dm.nv + 6 * dm.nbody; i++)
      if (s[i] != 0.0) { set_error("tokp_step_worlds_static: non-zero qfrc_applied/xfrc_applied unsupported"); return TOKP_ERR_UNSUPPORTED; }
  }
  unsigned char* base = nullptr;
  int rc = static_scratch_ptr(&base);
  if (rc) return rc;
  double* d_in = reinterpret_cast<double*>(base);
  double* d_out = d_in + (size_t)count * S;
  int* d_st = reinterpret_cast<int*>(d_out + (size_t)count * S);
  TOKP_CUDA(cudaMemcpy(d_in, snapshots_in, snap_bytes, cudaMemcpyHostToDevice));
  // the array keeps the last call's contents: outputs and status words start from zero
  TOKP_CUDA(cudaMemset(d_out, 0, snap_bytes + (size_t)count * sizeof(int)));
  int iters = solver_iterations < 0 ? dm.iterations : solver_iterations;
  double tol = solver_iterations < 0 ? dm.tolerance : solver_tolerance;
  if (!m->dev->lanes) m->dev->lanes = lanes_for_model(dm);
  int G = m->dev->lanes;
  for (int pass = 0; pass < (needs_redo_pass(dm) ? 2 : 1); pass++) {
    LaunchPlan lp;
    rc = plan_launch(dm, G, pass, &lp);
    if (rc) return rc;
    int blocks = (count + lp.worlds_per_block - 1) / lp.worlds_per_block;
#define LAUNCH(GG, NN)                                                                                                    \
  TOKP_CUDA(cudaFuncSetAttribute(k_step_worlds<GG, NN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lp.smem));      \
  k_step_worlds<GG, NN><<<blocks, lp.warps * 32, lp.smem>>>(dm, d_in, d_out, d_st, count, nsteps, iters, tol, S, lp.wsd, \
                                                        lp.ecap, pass, 0);
    TOKP_DISPATCH(G, dm.spec_nv, LAUNCH)
#undef LAUNCH
    g_launches++;
  }
  TOKP_CUDA(cudaGetLastError());
  TOKP_CUDA(cudaDeviceSynchronize());
  TOKP_CUDA(cudaMemcpy(snapshots_out, d_out, snap_bytes, cudaMemcpyDeviceToHost));
  if (status) TOKP_CUDA(cudaMemcpy(status, d_st, (size_t)count * sizeof(int), cudaMemcpyDeviceToHost));
  return TOKP_OK;
}

extern "C" int64_t tokp_launch_count(void) { return g_launches.load(); }
