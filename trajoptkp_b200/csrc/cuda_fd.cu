// Batched finite-difference dynamics derivatives: every (trajectory b, key-point t, perturbed column, +-eps)
// pair is one independent one-step world; one launch per iteration.
// Replaces Optimiser::ComputeDynamicsDerivativesAtKeypoints + WorkerComputeDerivatives
// (src/Optimiser/Optimiser.cpp:239-323) and the loop nest of Differentiator::DynamicsDerivatives
// (src/Differentiator/Differentiator.cpp:8-462).
//
// Stage reuse (the fork's mj_stepSkip, Differentiator.cpp:118,155 skip >= VEL for control columns, :253,276 skip >= POS
// for velocity columns, :357,377 none for positions): the worlds of a key-point that do not perturb qpos share the
// nominal world's position stage, the control worlds its velocity stage too, and both its factor of M + h D
// (mj_EulerSkip's skipfactor).  A launch therefore runs in two phases over two compacted work lists:
//   phase 0  nominal and position-perturbed worlds: full step; the nominal world stores what later stages read
//            (xipos, rootcom, cinert, cdof, M, its L'DL factor, J, D, B, K-terms | qfrc_passive, qfrc_bias, aref |
//            factor of M + h D) in the key-point's slot of the share buffer (HBM, L2-resident);
//   phase 1  control and velocity worlds: load instead of recompute, then run only the stages mj_stepSkip would run.
// The loaded values are the bits the world would have computed itself, so A / B are bit-identical to the
// every-world-runs-everything form (tests/test_gpu_determinism.py).
#include <cstdlib>

#include "step_dyn.cuh"
#include "traj.h"

extern "C" int tokp_fd_kernel_ms(tokp_traj* tr, double out[3]);
namespace tokp { int check_world_status(tokp_traj* tr, int column, const char* what); }

namespace tokp {

// world slot layout inside one key-point (W = 1 + 2m + 4n):
//   0                      nominal
//   1 + 2i + s             control i,  s = 0 (+eps) / 1 (-eps)
//   1 + 2m + 2i + s        velocity i
//   1 + 2m + 2n + 2i + s   position i
struct FdArgs {
  const double* snap;               // [B][T+1][SD]
  double* fd_out;                   // [B][T][W][nq+nv]
  int* status;                      // [B][T][W]: bit 30 = computed, low bits = step status
  const unsigned long long* mask;   // [B][T]
  const int* rows;                  // optional list of b*T+t rows (nullptr = all)
  int nrows;
  int B, T, SD, W, n, m;
  const int *qpos_adr, *dof_adr, *ctrl_adr;
  const double* ctrl_limits;
  int central;
  double eps;
  int iterations;
  double tolerance;
  int ecap, pass;                   // constraint-row capacity of this pass; pass 1 redoes overflowed worlds only
  int phase;                        // 0: nominal + position worlds (full step), 1: control + velocity worlds (stage reuse)
  int *list0, *list1;               // compacted world lists of pass 0 / pass 1 (gidx = row index * W + slot): phase-0
                                    // worlds fill a list from the front, phase-1 worlds from the back
  long long list_cap;               // entries per list
  int* counts;                      // [2 * pass + phase] list sizes, [4 + 2 * pass + phase] cursors
  int row_lo, row_hi;               // row-index range of this chunk (the share buffer holds row_hi - row_lo slots)
  double* share;                    // [row_hi - row_lo][share_stride]
  int share_stride;
  int share_on;                     // 0: every world runs the full step (phase 0), for A/B bit-identity checks
  int lane_mode;                    // 1: the control / velocity worlds run in k_fd_dep_lanes (one world per lane);
                                    // 2: ALL worlds do: the group kernel only runs the position stage of the nominal and
                                    //    position worlds (pos_only) and publishes one slot per world (slot = index in list0)
  int pos_only;
  int* rowslot;                     // [row_hi - row_lo] lane_mode 2: slot of the row's nominal world
  int blocksync;                    // Grp::bsync level of this launch (0 = warps run independently)
  int* wstatus;                     // [B][2] sticky per-trajectory warning bits (column 0 = FD worlds)
  int last;                         // no further pass re-runs the worlds that overflow in this one
};

// which perturbation a world slot carries: kind 0 nominal, 1 control, 2 velocity, 3 position
struct FdWorld {
  int row, slot, b, t, kind, col, sgn;
};

// decode world `gidx` (= row index * W + slot) and decide whether it is simulated at all
__device__ inline bool fd_decode(const DM& dm, const FdArgs& a, long long gidx, FdWorld& wd) {
  int rowi = (int)(gidx / a.W);
  wd.slot = (int)(gidx - (long long)rowi * a.W);
  wd.row = a.rows ? a.rows[rowi] : rowi;
  wd.b = wd.row / a.T; wd.t = wd.row - wd.b * a.T;
  wd.kind = 0; wd.col = 0; wd.sgn = 0;
  unsigned long long mask = a.mask[wd.row];
  if (mask == 0ull) return false;
  if (wd.slot > 0) {
    int q = wd.slot - 1;
    if (q < 2 * a.m) { wd.kind = 1; wd.col = q >> 1; wd.sgn = q & 1; }
    else if (q < 2 * a.m + 2 * a.n) { q -= 2 * a.m; wd.kind = 2; wd.col = q >> 1; wd.sgn = q & 1; }
    else { q -= 2 * a.m + 2 * a.n; wd.kind = 3; wd.col = q >> 1; wd.sgn = q & 1; }
    if (!((mask >> wd.col) & 1ull)) return false;
  }
  // control perturbations are clamp-aware (Differentiator.cpp:98-102,136-143); forward mode drops the minus
  // side unless the plus side is infeasible
  if (wd.kind == 1) {
    const double* s = a.snap + ((size_t)wd.b * (a.T + 1) + wd.t) * a.SD;
    double u = s[1 + dm.nq + 3 * dm.nv + a.ctrl_adr[wd.col]];
    bool fwd_ok = !(u + a.eps > a.ctrl_limits[2 * wd.col + 1]);
    bool back_wanted = a.central || !fwd_ok;
    bool back_ok = back_wanted && !(u - a.eps < a.ctrl_limits[2 * wd.col]);
    if ((wd.sgn == 0 && !fwd_ok) || (wd.sgn == 1 && !back_ok)) return false;
  } else if (wd.kind >= 2 && wd.sgn == 1 && !a.central) {
    return false;
  }
  return true;
}

// Work list of the launch: one thread per candidate world; the worlds that run are appended (warp-aggregated) to
// list0, all others get status 0.  Sparse key-point rows leave most of the B*T*W candidates empty; compaction fills
// every warp of k_fd_worlds with running worlds.
__global__ void k_fd_plan(DM dm, FdArgs a) {
  long long gidx = (long long)a.row_lo * a.W + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long total = (long long)a.row_hi * a.W;
  bool run = false;
  FdWorld wd;
  wd.kind = 0;
  if (gidx < total) {
    run = fd_decode(dm, a, gidx, wd);
    if (!run) a.status[(size_t)wd.row * a.W + wd.slot] = 0;
  }
  const bool dep = run && a.share_on && (wd.kind == 1 || wd.kind == 2);
  unsigned bal0 = __ballot_sync(0xffffffffu, run && !dep), bal1 = __ballot_sync(0xffffffffu, dep);
  int lane = threadIdx.x & 31;
  int base0 = 0, base1 = 0;
  if (lane == 0 && bal0) base0 = atomicAdd(&a.counts[0], __popc(bal0));
  if (lane == 0 && bal1) base1 = atomicAdd(&a.counts[1], __popc(bal1));
  base0 = __shfl_sync(0xffffffffu, base0, 0);
  base1 = __shfl_sync(0xffffffffu, base1, 0);
  unsigned below = (1u << lane) - 1u;
  if (run && !dep) {
    const int li = base0 + __popc(bal0 & below);
    a.list0[li] = (int)gidx;
    if (a.lane_mode == 2 && wd.kind == 0) {  // nominal world: its slot serves the whole key-point
      a.rowslot[(int)(gidx / a.W) - a.row_lo] = li;
      a.list1[atomicAdd(&a.counts[2], 1)] = li;
    }
  }
  if (dep) a.list0[a.list_cap - 1 - (base1 + __popc(bal1 & below))] = (int)gidx;
}

// share-buffer slot of one key-point (doubles): what the nominal world hands to the worlds that reuse its stages
struct ShareLayout {
  int hdr, M, LD, LDinv, J, D, aref, fpas, fbias, Bc, kterm, xipos, rootcom, cinert, cdof, LDe, Hinv, total;
  __host__ __device__ ShareLayout(const DM& m) {
    int nv = m.nv, nb = m.nbody, ne = m.maxefc, p = 0;
    auto take = [&](int n) { int r = p; p += n; return r; };
    hdr = take(4); M = take(nv * nv); LD = take(nv * nv); LDinv = take(nv); J = take(ne * nv); D = take(ne); aref = take(ne);
    fpas = take(nv); fbias = take(nv); Bc = take(ne); kterm = take(ne); xipos = take(3 * nb); rootcom = take(3 * nb);
    cinert = take(10 * nb); cdof = take(6 * nv); LDe = take(nv * nv); Hinv = take(nv);
    total = (p + 1) & ~1;
  }
};

template <int G>
__device__ __forceinline__ void cp_lanes(double* __restrict__ dst, const double* __restrict__ src, int n, int gl) {
  for (int i = gl; i < n; i += G) dst[i] = src[i];
}

// Persistent kernel: every warp fetches GPW worlds at a time from the compacted list (dynamic cursor), one group of
// G lanes per world.  pass 0 reads list0 and appends the worlds that overflow the fast constraint-row capacity to
// list1; pass 1 re-runs list1 with ecap = maxefc.
template <int G, int NV>
__global__ void __launch_bounds__(G == 16 ? 384 : 128, G == 16 ? 1 : 2) k_fd_worlds(DM dm, FdArgs a, int wsd) {
  extern __shared__ double smem[];
  constexpr int GPW = 32 / G;
  tk::Grp<G, NV> g;
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int slot_s = warp * GPW + lane / G;
  const int* list = a.pass == 0 ? a.list0 : a.list1;
  const int count = a.counts[2 * a.pass + a.phase];
  int* cursor = &a.counts[4 + 2 * a.pass + a.phase];
  int nq = dm.nq, nv = dm.nv;
  const ShareLayout L(dm);
  tk::WS w;
  tk::ws_init(w, smem + (size_t)slot_s * wsd, dm, a.ecap, G);
  __shared__ int s_base;
  for (;;) {
    int base = 0;
    if (a.blocksync) {
      // the whole block fetches one chunk of worlds and walks the stages of the step together
      __syncthreads();
      if (threadIdx.x == 0) s_base = atomicAdd(cursor, GPW * (int)(blockDim.x >> 5));
      __syncthreads();
      base = s_base;
      if (base >= count) break;
      base += warp * GPW;
      g.bsync = a.blocksync;
    } else {
      if (lane == 0) base = atomicAdd(cursor, GPW);
      base = __shfl_sync(0xffffffffu, base, 0);
      if (base >= count) break;
    }
    int li = base + lane / G;
    // block-synchronous mode: groups past the end of the list re-run the last world (every thread must reach the
    // block barriers inside the step) and drop the result
    const bool real = li < count;
    if (a.blocksync && !real) li = count - 1;
    bool run = li < count;
    g.wmask = __ballot_sync(0xffffffffu, run);
    long long gidx = 0;
    FdWorld wd;
    double time = 0.0;
    int st = 0;
    if (run) {
      gidx = list[a.phase == 0 ? li : a.list_cap - 1 - li];
      fd_decode(dm, a, gidx, wd);
      const double* s = a.snap + ((size_t)wd.b * (a.T + 1) + wd.t) * a.SD;
      time = s[0];
      for (int i = g.gl; i < nq; i += G) w.qpos[i] = s[1 + i];
      for (int i = g.gl; i < nv; i += G) {
        w.qvel[i] = s[1 + nq + i];
        w.qacc[i] = s[1 + nq + nv + i];
        w.warm[i] = s[1 + nq + 2 * nv + i];
      }
      for (int i = g.gl; i < dm.nu; i += G) w.ctrl[i] = s[1 + nq + 3 * nv + i];
      g.sync();
      if (g.gl == 0 && wd.kind != 0) {
        double e = wd.sgn ? -a.eps : a.eps;
        if (wd.kind == 1) {
          w.ctrl[a.ctrl_adr[wd.col]] += e;
        } else if (wd.kind == 2) {
          w.qvel[a.dof_adr[wd.col]] += e;
        } else {
          // mj_integratePos(qpos, e_dof, +-eps)
          int d = a.dof_adr[wd.col];
          int j = dm.dof_jntid[d];
          int pa = dm.jnt_qposadr[j], va = dm.jnt_dofadr[j];
          if (dm.jnt_type[j] == tk::JNT_FREE) {
            int kk = d - va;
            if (kk < 3) w.qpos[pa + kk] += e * 1.0;
            else {
              double vel[3] = {0.0, 0.0, 0.0};
              vel[kk - 3] = 1.0;
              tk::quat_integrate(&w.qpos[pa + 3], vel, e);
            }
          } else {
            w.qpos[pa] += e * 1.0;
          }
        }
      }
      g.sync();
      double* sh = a.share + (size_t)(a.pos_only ? li : (int)(gidx / a.W) - a.row_lo) * a.share_stride;
      w.work = 0;
      if (a.pos_only) {
        // lane_mode 2: position stage only (mj_fwdPosition) + the factor of M + h D; everything after it runs one world
        // per lane in k_fd_dep_lanes, which reads this world's slot
        tk::step_position(dm, w, g, st);
        if (real) {
          const int ne = w.nefc;
          if (g.gl == 0) { sh[L.hdr] = (double)ne; sh[L.hdr + 2] = 0.0; }
          cp_lanes<G>(sh + L.M, w.M, nv * nv, g.gl); cp_lanes<G>(sh + L.LD, w.LD, nv * nv, g.gl);
          cp_lanes<G>(sh + L.LDinv, w.LDinv, nv, g.gl); cp_lanes<G>(sh + L.J, w.J, ne * nv, g.gl);
          cp_lanes<G>(sh + L.D, w.efc_D, ne, g.gl);
          cp_lanes<G>(sh + L.Bc, w.efc_B, ne, g.gl); cp_lanes<G>(sh + L.kterm, w.efc_kterm, ne, g.gl);
          cp_lanes<G>(sh + L.cinert, w.cinert, 10 * dm.nbody, g.gl); cp_lanes<G>(sh + L.cdof, w.cdof, 6 * nv, g.gl);
        }
        g.phase();
        tk::euler_factor(dm, w, g);
        g.sync();
        if (real) {
          cp_lanes<G>(sh + L.LDe, w.LD, nv * nv, g.gl); cp_lanes<G>(sh + L.Hinv, w.Hinv, nv, g.gl);
          int bits = 0;
          for (int k = 0; k < 6; k++) if (g.any((st >> k) & 1)) bits |= (1 << k);
          if (g.gl == 0) sh[L.hdr + 1] = (double)bits;
        }
        g.sync();
      } else if (a.phase == 0) {
        // full step (mj_step / mj_stepSkip(mjSTAGE_NONE)); the nominal world publishes its stages
        tk::step_position(dm, w, g, st);
        tk::step_velocity(dm, w, g);
        const bool pub = wd.kind == 0 && real;
        if (pub) {
          const int ne = w.nefc;
          if (g.gl == 0) { sh[L.hdr] = (double)ne; sh[L.hdr + 1] = 0.0; }
          cp_lanes<G>(sh + L.M, w.M, nv * nv, g.gl); cp_lanes<G>(sh + L.LD, w.LD, nv * nv, g.gl);
          cp_lanes<G>(sh + L.LDinv, w.LDinv, nv, g.gl); cp_lanes<G>(sh + L.J, w.J, ne * nv, g.gl);
          cp_lanes<G>(sh + L.D, w.efc_D, ne, g.gl); cp_lanes<G>(sh + L.aref, w.efc_aref, ne, g.gl);
          cp_lanes<G>(sh + L.fpas, w.frc_passive, nv, g.gl); cp_lanes<G>(sh + L.fbias, w.frc_bias, nv, g.gl);
          cp_lanes<G>(sh + L.Bc, w.efc_B, ne, g.gl); cp_lanes<G>(sh + L.kterm, w.efc_kterm, ne, g.gl);
          cp_lanes<G>(sh + L.xipos, w.xipos, 3 * dm.nbody, g.gl); cp_lanes<G>(sh + L.rootcom, w.rootcom, 3 * dm.nbody, g.gl);
          cp_lanes<G>(sh + L.cinert, w.cinert, 10 * dm.nbody, g.gl); cp_lanes<G>(sh + L.cdof, w.cdof, 6 * nv, g.gl);
        }
        g.phase();
        tk::fwd_constraint(dm, w, g, a.iterations, a.tolerance);
        g.phase();
        tk::check_qacc(dm, w, g, st);
        tk::euler_factor(dm, w, g);
        if (pub) {
          g.sync();
          cp_lanes<G>(sh + L.LDe, w.LD, nv * nv, g.gl); cp_lanes<G>(sh + L.Hinv, w.Hinv, nv, g.gl);
          // status bits of the position / velocity stage travel with the slot (overflow: the dependants are redone too)
          int bits = 0;
          for (int k = 0; k < 6; k++) if (g.any((st >> k) & 1)) bits |= (1 << k);
          if (g.gl == 0) {
            sh[L.hdr + 1] = (double)(bits & (tk::ST_OVERFLOW | tk::ST_UNSUPPORTED | tk::ST_BAD_QPOS));
            sh[L.hdr + 2] = (double)w.work;  // line-search evaluations of the nominal world: its neighbours will need as many
          }
        }
        tk::euler_advance(dm, w, g, time);
      } else {
        // mj_stepSkip(mjSTAGE_POS) for velocity worlds, mj_stepSkip(mjSTAGE_VEL) for control worlds
        const int ne = (int)sh[L.hdr];
        st |= (int)sh[L.hdr + 1];
        w.nefc = ne;
        for (int i = g.gl; i < nq; i += G) if (tk::bad_value(w.qpos[i])) st |= tk::ST_BAD_QPOS;
        for (int i = g.gl; i < nv; i += G) if (tk::bad_value(w.qvel[i])) st |= tk::ST_BAD_QVEL;
        cp_lanes<G>(w.M, sh + L.M, nv * nv, g.gl); cp_lanes<G>(w.LD, sh + L.LD, nv * nv, g.gl);
        cp_lanes<G>(w.LDinv, sh + L.LDinv, nv, g.gl); cp_lanes<G>(w.J, sh + L.J, ne * nv, g.gl);
        cp_lanes<G>(w.efc_D, sh + L.D, ne, g.gl);
        if (wd.kind == 1) {
          cp_lanes<G>(w.efc_aref, sh + L.aref, ne, g.gl);
          cp_lanes<G>(w.frc_passive, sh + L.fpas, nv, g.gl); cp_lanes<G>(w.frc_bias, sh + L.fbias, nv, g.gl);
        } else {
          cp_lanes<G>(w.efc_B, sh + L.Bc, ne, g.gl); cp_lanes<G>(w.efc_kterm, sh + L.kterm, ne, g.gl);
          cp_lanes<G>(w.xipos, sh + L.xipos, 3 * dm.nbody, g.gl); cp_lanes<G>(w.rootcom, sh + L.rootcom, 3 * dm.nbody, g.gl);
          cp_lanes<G>(w.cinert, sh + L.cinert, 10 * dm.nbody, g.gl); cp_lanes<G>(w.cdof, sh + L.cdof, 6 * nv, g.gl);
        }
        g.sync();
        g.phase();
        if (wd.kind == 2) {
          tk::com_vel(dm, w, g);
          tk::passive_and_bias(dm, w, g);
        }
        g.phase();
        if (wd.kind == 2) tk::reference_constraint(dm, w, g);
        tk::actuation_acceleration(dm, w, g);
        g.phase();
        tk::fwd_constraint(dm, w, g, a.iterations, a.tolerance);
        g.phase();
        tk::check_qacc(dm, w, g, st);
        if (dm.any_damping) {  // mj_EulerSkip(skipfactor = 1): the nominal world's factor of M + h D
          cp_lanes<G>(w.LD, sh + L.LDe, nv * nv, g.gl); cp_lanes<G>(w.Hinv, sh + L.Hinv, nv, g.gl);
          g.sync();
        }
        tk::euler_advance(dm, w, g, time);
      }
    }
    if (run && real && !a.pos_only) {
      double* o = a.fd_out + ((size_t)wd.row * a.W + wd.slot) * (nq + nv);
      for (int i = g.gl; i < nq; i += G) o[i] = w.qpos[i];
      for (int i = g.gl; i < nv; i += G) o[nq + i] = w.qvel[i];
      int bits = 0;
      for (int k = 0; k < 6; k++) if (g.any((st >> k) & 1)) bits |= (1 << k);
      if (g.gl == 0) {
        a.status[(size_t)wd.row * a.W + wd.slot] = bits | (1 << 30);
        if (a.pass == 0 && (bits & tk::ST_OVERFLOW)) {
          int k = atomicAdd(&a.counts[2 + a.phase], 1);
          a.list1[a.phase == 0 ? k : a.list_cap - 1 - k] = (int)gidx;
        }
        int keep = a.last ? bits : (bits & ~tk::ST_OVERFLOW);
        if (keep) atomicOr(&a.wstatus[2 * wd.b], keep);
      }
      g.sync();
    }
    __syncwarp();
  }
}

}  // namespace tokp
#include "fd_lanes.cuh"
namespace tokp {

// (q2 - q1)/h on the configuration manifold for one dof (mj_differentiatePos restricted to a dof)
__device__ inline double diff_pos_dof(const DM& dm, const double* q1, const double* q2, double h, int d) {
  int j = dm.dof_jntid[d];
  int pa = dm.jnt_qposadr[j], va = dm.jnt_dofadr[j];
  if (dm.jnt_type[j] == tk::JNT_FREE) {
    int kk = d - va;
    if (kk < 3) return (q2[pa + kk] - q1[pa + kk]) / h;
    double v[3];
    tk::sub_quat(v, q1 + pa + 3, q2 + pa + 3);
    return v[kk - 3] / h;
  }
  return (q2[pa] - q1[pa]) / h;
}

// one thread per (row, kind, column, state row): differencing + scatter (Differentiator.cpp:167-222,286-321,386-457)
__global__ void k_fd_assemble(DM dm, FdArgs a, double* __restrict__ A, double* __restrict__ Bm) {
  int nx = 2 * a.n;
  long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long per_row = (long long)(a.m + 2 * a.n) * nx;
  long long total = (long long)(a.row_hi - a.row_lo) * per_row;
  if (gid >= total) return;
  int rowi = a.row_lo + (int)(gid / per_row);
  int rem = (int)(gid - (long long)(rowi - a.row_lo) * per_row);
  int c = rem / nx, j = rem - c * nx;
  int row = a.rows ? a.rows[rowi] : rowi;
  int kind, col;
  if (c < a.m) { kind = 1; col = c; }
  else if (c < a.m + a.n) { kind = 2; col = c - a.m; }
  else { kind = 3; col = c - a.m - a.n; }
  unsigned long long mask = a.mask[row];
  if (!((mask >> col) & 1ull)) return;
  int nq = dm.nq, nv = dm.nv, S2 = nq + nv;
  int base = (kind == 1) ? 1 : (kind == 2 ? 1 + 2 * a.m : 1 + 2 * a.m + 2 * a.n);
  int sp = base + 2 * col, sm = sp + 1;
  const int* st = a.status + (size_t)row * a.W;
  bool vp = (st[sp] >> 30) & 1, vm = (st[sm] >> 30) & 1;
  if (!vp && !vm) return;
  const double* out0 = a.fd_out + (size_t)row * a.W * S2;
  const double *qa, *qb;
  double h;
  if (vp && vm) { qa = out0 + (size_t)sm * S2; qb = out0 + (size_t)sp * S2; h = 2 * a.eps; }
  else if (vp) { qa = out0; qb = out0 + (size_t)sp * S2; h = a.eps; }
  else { qa = out0 + (size_t)sm * S2; qb = out0; h = a.eps; }
  double val;
  if (j < a.n) val = diff_pos_dof(dm, qa, qb, h, a.dof_adr[j]);
  else { int d = a.dof_adr[j - a.n]; val = (qb[nq + d] - qa[nq + d]) / h; }
  if (kind == 1) Bm[(size_t)row * nx * a.m + (size_t)col * nx + j] = val;
  else if (kind == 2) A[(size_t)row * nx * nx + (size_t)(col + a.n) * nx + j] = val;
  else A[(size_t)row * nx * nx + (size_t)col * nx + j] = val;
}

// launch helper shared by tokp_dynamics_derivs / _cols / iterative_error
int launch_fd(tokp_traj* tr, const unsigned long long* mask, const int* rows, int nrows, int central, double eps) {
  const DM& dm = tr->dm;
  FdArgs a;
  a.snap = tr->snap; a.fd_out = tr->fd_out; a.status = tr->status; a.mask = mask; a.rows = rows; a.nrows = nrows;
  a.B = tr->B; a.T = tr->T; a.SD = tr->SD; a.W = tr->W; a.n = tr->n; a.m = tr->m;
  a.qpos_adr = tr->task.qpos_adr; a.dof_adr = tr->task.dof_adr; a.ctrl_adr = tr->task.ctrl_adr;
  a.ctrl_limits = tr->task.ctrl_limits;
  a.central = central; a.eps = eps;
  a.iterations = 5; a.tolerance = 0.0;  // MuJoCoHelper::InitModelForFiniteDifferencing (MuJoCoHelper.cpp:925-931)
  int G = tr->G;
  long long worlds = (long long)nrows * tr->W;
  if (worlds < 1) return TOKP_OK;
  if (worlds >= (1ll << 31)) { set_error("too many FD worlds for one launch"); return TOKP_ERR_ARG; }
  if (tr->fd_timed) {  // fold the previous launch's duration into the running sum before reusing the events
    double tmp[3];
    tokp_fd_kernel_ms(tr, tmp);
  }
  // lane-per-world kernel for the worlds of small hinge / slide models with an nv-specialised instantiation:
  // fd_lanes 1 = the stage-reusing (control / velocity) worlds, 2 = every world after a position-stage-only group kernel
  const int lane_mode = (tr->fd_stage_sharing && tr->fd_lanes && dm.lane_ok && 2 * (tr->m + tr->n) <= 64) ? (tr->fd_lanes >= 2 ? 2 : 1) : 0;
  // share buffer: one slot per key-point row of a chunk (lane_mode 2: one per nominal / position world, 1 + 2n per row),
  // at most ~1 GiB (fd_share_gib, default 16 GiB, in lane_mode 2: 180 GB of HBM make that cheap, and one chunk has one set
  // of kernel tails); larger launches run in row chunks
  const ShareLayout L(dm);
  const int slots_per_row = lane_mode == 2 ? 1 + 2 * tr->n : 1;
  if (tr->fd_share && tr->fd_share_slots_per_row != slots_per_row) {
    TOKP_CUDA(cudaStreamSynchronize(tr->stream));
    cudaFree(tr->fd_share);
    tr->fd_share = nullptr;
  }
  if (!tr->fd_share) {
    long long want = (long long)tr->B * tr->T;
    long long cap = ((long long)(lane_mode == 2 ? tr->fd_share_gib : 1) << 30) / ((long long)L.total * 8 * slots_per_row);
    if (cap < 1) cap = 1;
    tr->fd_share_rows = (int)(want < cap ? want : cap);
    tr->fd_share_slots_per_row = slots_per_row;
    size_t share_bytes = (size_t)tr->fd_share_rows * slots_per_row * L.total * 8;
    TOKP_CUDA(cudaMalloc(&tr->fd_share, share_bytes + (size_t)tr->fd_share_rows * sizeof(int)));
    tr->fd_rowslot = reinterpret_cast<int*>(reinterpret_cast<unsigned char*>(tr->fd_share) + share_bytes);
  }
  a.list0 = tr->fd_list0; a.list1 = tr->fd_list1; a.counts = tr->fd_counts; a.wstatus = tr->wstatus;
  a.list_cap = (long long)tr->B * tr->T * tr->W;
  a.share = tr->fd_share; a.share_stride = L.total; a.share_on = tr->fd_stage_sharing;
  a.lane_mode = lane_mode; a.pos_only = 0; a.rowslot = tr->fd_rowslot;
  tr->fd_last_lane_mode = lane_mode;
  // block-synchronous walk of the step (see Grp::bsync): the warps of a block stay in the same region of the step
  // code, which is several times the instruction cache.  TOKP_FD_BLOCKSYNC=0 lets the warps run independently.
  a.blocksync = tr->fd_blocksync;
  cudaEventRecord(tr->ev2, tr->stream);
  // even chunks: every chunk pays the same kernel tails, a short remainder chunk would pay them for little work
  const int nchunks = (nrows + tr->fd_share_rows - 1) / tr->fd_share_rows;
  const int chunk_rows = nchunks > 0 ? (nrows + nchunks - 1) / nchunks : 1;
  for (int row_lo = 0; row_lo < nrows; row_lo += chunk_rows) {
    a.row_lo = row_lo;
    a.row_hi = row_lo + chunk_rows < nrows ? row_lo + chunk_rows : nrows;
    long long cw = (long long)(a.row_hi - a.row_lo) * tr->W;
    a.ecap = 0; a.pass = 0; a.phase = 0; a.last = 0;
    TOKP_CUDA(cudaMemsetAsync(tr->fd_counts, 0, 8 * sizeof(int), tr->stream));
    k_fd_plan<<<(int)((cw + 255) / 256), 256, 0, tr->stream>>>(dm, a);
    g_launches++;
    tr->counters[2] += 1;

    for (int pass = 0; pass < (a.lane_mode == 2 ? 1 : (needs_redo_pass(dm) ? 2 : 1)); pass++) {
      LaunchPlan lp;
      // lane_mode 2: the group kernel runs the position stage only, once, at the full constraint-row capacity
      int rc = plan_launch(dm, G, a.lane_mode == 2 ? 1 : pass, &lp);
      if (rc) return rc;
      a.ecap = lp.ecap; a.pass = pass; a.last = (pass == 1 || !needs_redo_pass(dm) || a.lane_mode == 2) ? 1 : 0;
      a.pos_only = a.lane_mode == 2 ? 1 : 0;
      // persistent grid: as many blocks as stay resident, never more than the worlds need
      long long need = (cw + lp.worlds_per_block - 1) / lp.worlds_per_block;
      int blocks = (int)(need < (long long)lp.resident_blocks ? need : (long long)lp.resident_blocks);
      for (int phase = 0; phase < (a.lane_mode ? 1 : 2); phase++) {
        a.phase = phase;
#define LAUNCH(GG, NN)                                                                                               \
  TOKP_CUDA(cudaFuncSetAttribute(k_fd_worlds<GG, NN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lp.smem));   \
  k_fd_worlds<GG, NN><<<blocks, lp.warps * 32, lp.smem, tr->stream>>>(dm, a, lp.wsd);
        TOKP_DISPATCH(G, dm.spec_nv, LAUNCH)
#undef LAUNCH
        g_launches++;
        tr->counters[2] += 1;
      }
    }
    if (a.lane_mode) {
      // after every pass of the group kernel the slots are final.  lane_mode 2 first runs the nominal worlds (they
      // leave their work count and velocity stage in the slot); then the remaining worlds are binned by constraint rows
      // and the nominal world's work and run in the light bins (small workspace, more warps per SM) and the heavy ones.
      LaneArgs la;
      la.counts = tr->fd_counts + 8;  // [0, kLaneBins) sizes | [kLaneBins, 2 kLaneBins) fill counters | 2 cursors
      la.list = tr->fd_list2;
      la.necap = 0; la.heavy = 0; la.full = a.lane_mode == 2 ? 1 : 0; la.stage = 1;
      a.pass = 0; a.phase = 1; a.last = 1; a.pos_only = 0;
      auto launch_lanes = [&](int necap) -> int {
        la.necap = necap;
        const LanePlan lpn(dm, la.necap);
        size_t lsmem = (size_t)lpn.total * 8;
        int per_sm = (int)((227 * 1024) / (lsmem + 1024));
        if (per_sm < 1) { set_error("lane workspace does not fit shared memory"); return TOKP_ERR_UNSUPPORTED; }
        long long lneed = (cw + 31) / 32;
        int lblocks = (int)(lneed < (long long)per_sm * tr->sm_count ? lneed : (long long)per_sm * tr->sm_count);
#define LAUNCH_LANES(NN)                                                                                              \
  TOKP_CUDA(cudaFuncSetAttribute(k_fd_dep_lanes<NN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lsmem));       \
  k_fd_dep_lanes<NN><<<lblocks, 32, lsmem, tr->stream>>>(dm, a, la);
        if (dm.spec_nv == 9) { LAUNCH_LANES(9) } else { LAUNCH_LANES(2) }
#undef LAUNCH_LANES
        g_launches++;
        tr->counters[2] += 1;
        return TOKP_OK;
      };
      for (int stage = la.full ? 0 : 1; stage < 2; stage++) {
        la.stage = stage;
        TOKP_CUDA(cudaMemsetAsync(la.counts, 0, (2 * kLaneBins + 2) * sizeof(int), tr->stream));
        k_fd_bin<<<(int)((cw + 255) / 256), 256, 0, tr->stream>>>(dm, a, la, nullptr, 0);
        k_fd_bin<<<(int)((cw + 255) / 256), 256, 0, tr->stream>>>(dm, a, la, tr->fd_list2, 1);
        g_launches += 2;
        tr->counters[2] += 2;
        for (int heavy = 0; heavy < 2; heavy++) {
          la.heavy = heavy;
          if (heavy && dm.maxefc <= kLaneLightRows) break;
          int rc = launch_lanes(heavy ? dm.maxefc : (dm.maxefc < kLaneLightRows ? dm.maxefc : kLaneLightRows));
          if (rc) return rc;
        }
      }
    }
    long long tot = (long long)(a.row_hi - a.row_lo) * (tr->m + 2 * tr->n) * tr->nx;
    k_fd_assemble<<<(int)((tot + 255) / 256), 256, 0, tr->stream>>>(dm, a, tr->A, tr->Bm);
    g_launches++;
    tr->counters[2] += 1;
  }
  cudaEventRecord(tr->ev3, tr->stream);
  tr->fd_timed = true;
  TOKP_CUDA(cudaGetLastError());
  return TOKP_OK;
}

// contact-rich batches (push scenes: 30-120 rows) would send most worlds through the low-occupancy redo pass: when
// more than 5 % of a launch overflowed, the next launches of this batch start from twice the capacity
void adapt_fast_capacity(tokp_traj* tr, long long worlds, long long overflowed) {
  if (worlds > 0 && overflowed * 20 > worlds) {
    int cap = fast_ecap(tr->dm);
    if (cap < tr->dm.maxefc) tr->dm.fast_ecap = cap * 2 < tr->dm.maxefc ? cap * 2 : tr->dm.maxefc;
  }
}

// Key-point statistics of one trajectory per block, without copying the masks to the host: FD worlds and columns of the
// current masks, and the sum over the dofs of the percentage of rows that are key-points of that dof - with the
// velocity_change quirk of the last row (the scan's own hits, then every dof once more; KeyPointGenerator.cpp:724-727),
// per dof ((double)count / T) * 100 summed in dof order exactly as read_keypoints does on the host.
__global__ void k_kp_stats(const unsigned long long* __restrict__ mask, const unsigned long long* __restrict__ last,
                           const int* __restrict__ flags, double* __restrict__ out, int T, int n, int m) {
  __shared__ double pct[64];
  const int b = blockIdx.x, lane = threadIdx.x;
  const unsigned long long* mk = mask + (size_t)b * T;
  const bool quirk = (flags[b] & 1) != 0;
  const unsigned long long cmask = m >= 64 ? ~0ull : ((1ull << m) - 1);
  long long worlds = 0, cols = 0;
  for (int t = lane; t < T; t += 32) {
    const unsigned long long v = mk[t];
    if (!v) continue;
    const int c = __popcll(v);
    worlds += 1 + 2 * __popcll(v & cmask) + 4 * c;
    cols += c;
  }
  for (int off = 16; off > 0; off >>= 1) {
    worlds += __shfl_xor_sync(0xffffffffu, worlds, off);
    cols += __shfl_xor_sync(0xffffffffu, cols, off);
  }
  for (int j = lane; j < n; j += 32) {
    int cnt = 0;
    for (int t = 0; t < T; t++) {
      const unsigned long long v = (t == T - 1 && quirk) ? last[b] : mk[t];
      cnt += (int)((v >> j) & 1ull);
    }
    if (quirk) cnt++;
    pct[j] = ((double)cnt / (double)T) * 100;
  }
  __syncwarp();
  if (lane == 0) {
    double s = 0;
    for (int j = 0; j < n; j++) s += pct[j];
    out[3 * b] = s; out[3 * b + 1] = (double)worlds; out[3 * b + 2] = (double)cols;
  }
}

// the statistics, the FD list counters and the status words travel to pinned host memory behind the work already on
// the stream; nothing waits here
int kp_stats_enqueue(tokp_traj* tr) {
  const int B = tr->B;
  k_kp_stats<<<B, 32, 0, tr->stream>>>(tr->kp_mask, tr->kp_last, tr->kp_flags, tr->kp_stats, tr->T, tr->n, tr->m);
  g_launches++;
  TOKP_CUDA(cudaGetLastError());
  TOKP_CUDA(cudaMemcpyAsync(tr->h_kp_stats, tr->kp_stats, (size_t)B * 3 * 8, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaMemcpyAsync(tr->h_fd_counts, tr->fd_counts, 8 * sizeof(int), cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaMemcpyAsync(tr->h_wstatus, tr->wstatus, (size_t)B * 2 * sizeof(int), cudaMemcpyDeviceToHost, tr->stream));
  return TOKP_OK;
}

// after the stream has been synchronised: counters (unless the caller keeps its own, as the iterative_error level loop
// does) and the status words of the derivative phase
int kp_stats_finish(tokp_traj* tr, bool set_counters, const char* what) {
  if (set_counters) {
    double worlds = 0, cols = 0;
    for (int b = 0; b < tr->B; b++) { worlds += tr->h_kp_stats[3 * b + 1]; cols += tr->h_kp_stats[3 * b + 2]; }
    tr->counters[0] = (long long)worlds;
    // worlds that overflowed the fast constraint-row capacity and were re-run (of the last row chunk)
    const int* hc = tr->h_fd_counts;
    if (tr->fd_last_lane_mode == 2) {  // no fast pass: counts[2] is the number of nominal worlds there
      tr->counters[3] = 0;
    } else {
      tr->counters[3] = hc[2] + hc[3];
      adapt_fast_capacity(tr, hc[0] + hc[1], hc[2] + hc[3]);
    }
    tr->counters[1] = (int64_t)(cols / tr->n * 1e6);
  }
  if (!what) return TOKP_OK;
  for (int b = 0; b < tr->B; b++) {
    const int bits = tr->h_wstatus[2 * b];
    if (bits & 32) { set_error(std::string(what) + ": a world hit an unsupported collision pair or contact dimension"); return TOKP_ERR_UNSUPPORTED; }
    if (bits & 8) {
      set_error(std::string(what) + ": constraint rows of an FD world exceed TOKP_EFC_MAX, contacts were dropped (trajectory " +
                std::to_string(b) + ")");
      return TOKP_ERR_STATE;
    }
  }
  return TOKP_OK;
}

// counters: worlds actually simulated and Jacobians (|C|/n per row) for the current key-point masks
int update_fd_counters(tokp_traj* tr) {
  int rc = kp_stats_enqueue(tr);
  if (rc) return rc;
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  return kp_stats_finish(tr, true, nullptr);
}

}  // namespace tokp

using namespace tokp;

extern "C" int tokp_dynamics_derivs(tokp_traj* tr, int central_diff, double eps) {
  if (!tr || !(eps > 0)) { set_error("tokp_dynamics_derivs: bad argument"); return TOKP_ERR_ARG; }
  cudaEventRecord(tr->ev0, tr->stream);
  int rc = launch_fd(tr, tr->kp_mask, nullptr, tr->B * tr->T, central_diff, eps);
  if (rc) return rc;
  cudaEventRecord(tr->ev1, tr->stream);
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, tr->ev0, tr->ev1);
  tr->timers[0] = ms;
  int rc2 = update_fd_counters(tr);
  if (rc2) return rc2;
  return check_world_status(tr, 0, "tokp_dynamics_derivs");
}

extern "C" int tokp_dynamics_derivs_cols(tokp_traj* tr, int b, int t, const int* cols, int ncols, int central_diff,
                                         double eps, double* A, int lda, double* B, int ldb) {
  if (!tr || !cols || !A || !B || b < 0 || b >= tr->B || t < 0 || t >= tr->T || ncols < 0 || lda < tr->nx || ldb < tr->nx) {
    set_error("tokp_dynamics_derivs_cols: bad argument");
    return TOKP_ERR_ARG;
  }
  unsigned long long mk = 0;
  for (int i = 0; i < ncols; i++) {
    if (cols[i] < 0 || cols[i] >= tr->n) { set_error("tokp_dynamics_derivs_cols: column out of range"); return TOKP_ERR_ARG; }
    mk |= 1ull << cols[i];
  }
  int row = b * tr->T + t;
  // persistent scratch of the trajectory (allocated once, on the first call): an all-zero [B*T] mask in which only this
  // call's row is set for the duration of the launch, and the one-entry row list
  if (!tr->cols_scratch) {
    TOKP_CUDA(cudaMalloc(&tr->cols_scratch, (size_t)tr->B * tr->T * 8 + 8));
    TOKP_CUDA(cudaMemsetAsync(tr->cols_scratch, 0, (size_t)tr->B * tr->T * 8 + 8, tr->stream));
  }
  unsigned long long* d_mask = static_cast<unsigned long long*>(tr->cols_scratch);
  int* d_row = reinterpret_cast<int*>(d_mask + (size_t)tr->B * tr->T);
  TOKP_CUDA(cudaMemcpyAsync(d_mask + row, &mk, 8, cudaMemcpyHostToDevice, tr->stream));
  TOKP_CUDA(cudaMemcpyAsync(d_row, &row, 4, cudaMemcpyHostToDevice, tr->stream));
  int rc = launch_fd(tr, d_mask, d_row, 1, central_diff, eps);
  cudaMemsetAsync(d_mask + row, 0, 8, tr->stream);  // leave the mask all-zero again, whatever happened
  if (rc) return rc;
  int nx = tr->nx, m = tr->m, n = tr->n;
  std::vector<double> hA((size_t)nx * nx), hB((size_t)nx * (m > 0 ? m : 1));
  TOKP_CUDA(cudaMemcpyAsync(hA.data(), tr->A + (size_t)row * nx * nx, hA.size() * 8, cudaMemcpyDeviceToHost, tr->stream));
  if (m > 0) TOKP_CUDA(cudaMemcpyAsync(hB.data(), tr->Bm + (size_t)row * nx * m, (size_t)nx * m * 8, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  for (int i = 0; i < ncols; i++) {
    int c = cols[i];
    for (int j = 0; j < nx; j++) {
      A[(size_t)c * lda + j] = hA[(size_t)c * nx + j];
      A[(size_t)(c + n) * lda + j] = hA[(size_t)(c + n) * nx + j];
      if (c < m) B[(size_t)c * ldb + j] = hB[(size_t)c * nx + j];
    }
  }
  return TOKP_OK;
}
