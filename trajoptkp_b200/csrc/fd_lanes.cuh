// Lane-per-world kernel for the FD worlds that reuse the nominal world's stages (mj_stepSkip(mjSTAGE_POS) for
// velocity columns, mj_stepSkip(mjSTAGE_VEL) for control columns; Differentiator.cpp:109-163,242-283).
//
// Every lane runs ONE control / velocity world from start to end, sequentially, like a host thread would; a warp takes
// 32 consecutive worlds of the compacted list, i.e. the stage-reusing worlds of a handful of neighbouring key-points
// (sparse key-point rows - velocity_change, adaptive_jerk - have only 2-8 such worlds each).  What makes this work:
//   * everything the nominal world published (M, its L'DL factor, J, D, B, K-terms, cinert, cdof, the factor of M + h D)
//     is read-only here and shared by all worlds of a key-point: lanes read it straight from the key-point's slot of the
//     share buffer through L1 (lanes of the same key-point hit the same address);
//   * the private state of a world after the position stage is small (a dozen nv-vectors, two nefc-vectors, one
//     nv x nv Cholesky factor): it lives in shared memory as [index][lane] (bank-conflict free) or in registers;
//   * the worlds of a key-point differ by 1e-6 perturbations and neighbouring key-points are similar, so the lanes take
//     the same branches (same active rows, same line-search path) most of the time; where they do not the hardware
//     serialises them;
//   * no lane repeats scalar work of another lane, there are no shuffles and no group barriers, and the code is small
//     enough for the instruction cache, so the warps of an SM need no block barrier to stay together.
// The arithmetic is the sequential algorithm (same order as a CPU implementation would use), not the butterfly
// reductions of the group kernels, so results differ from those by rounding only.
// Included by cuda_fd.cu (uses FdArgs, FdWorld, fd_decode, ShareLayout).
#pragma once

namespace tokp {

// shared-memory plan of one warp (doubles)
struct LanePlan {
  int vec;      // 9 per-lane nv-vectors [9][nv][32]
  int scratch;  // per-lane union: velocity stage (cvel | cacc : 12 nb; cfrc borrows the solver's vectors when they are
                // large enough, else follows) / solver (jar | jv | L : 2 ne + ntri)
  int cfrc_in_vec;
  int total;
  __host__ __device__ LanePlan(const DM& m, int ne) {
    int nv = m.nv, nb = m.nbody;
    vec = 0;
    scratch = 9 * nv * 32;
    int ntri = nv * (nv + 1) / 2;
    cfrc_in_vec = 6 * nv >= 6 * nb;  // v_qacc .. v_fc (6 vectors) are dead during the velocity stage
    int a = (cfrc_in_vec ? 12 : 18) * nb, b = 2 * ne + ntri;
    total = scratch + (a > b ? a : b) * 32;
  }
};

// The stage-reusing worlds are sorted before they run, by how much solver work their key-point's nominal world needed
// (its number of line-search evaluations: a world that differs from it by a 1e-6 perturbation needs the same, almost
// always): a warp then holds worlds that run the same number of Newton iterations and evaluations, instead of waiting
// for its slowest lane - free-flight worlds (no solver), resting contacts (warm start already optimal: one iteration)
// and contact transitions (all five iterations) differ by an order of magnitude.  The major key is the row class:
// key-points with at most kLaneLightRows constraint rows run with a small workspace (more warps per SM).
// The minor key separates the control worlds (no velocity stage of their own) from the others.
constexpr int kLaneWorkBins = 48, kLaneHalf = 2 * kLaneWorkBins, kLaneBins = 2 * kLaneHalf;
constexpr int kLaneLightRows = 24;
__host__ __device__ inline int lane_bin(int ne, int work, int control) {
  int wbin = ne == 0 ? 0 : (work >= kLaneWorkBins - 1 ? kLaneWorkBins - 1 : 1 + work);
  if (wbin >= kLaneWorkBins) wbin = kLaneWorkBins - 1;
  return (ne > kLaneLightRows ? kLaneHalf : 0) + 2 * wbin + (control ? 1 : 0);
}

struct LaneArgs {
  int necap;        // constraint-row capacity of the warp workspace
  int heavy;        // 0: bins [0, kLaneHalf), 1: bins [kLaneHalf, kLaneBins)
  int full;         // 1: every world of the launch runs here after a position-stage-only group kernel (FdArgs::lane_mode 2);
                    //    list entries are then e >= 0: index into the front of list0 (nominal / position world, own slot e),
                    //    e < 0: stage-reusing world -1 - e of the back of list0 (slot of its key-point's nominal world)
  int stage;        // full mode: 0 = the nominal worlds (FdArgs::list1 holds their list0 indices; sorted by their own
                    // constraint-row count, the only thing known about them yet), 1 = all others
  const int* list;  // worlds sorted by bin (k_fd_bin_scatter)
  int* counts;      // [0, kLaneBins) bin sizes, [kLaneBins, 2 kLaneBins) fill counters of the scatter, then 2 cursors
};

// bin sizes of the stage-reusing worlds of this chunk (scatter = 0), then their placement (scatter = 1)
__global__ void k_fd_bin(DM dm, FdArgs a, LaneArgs la, int* __restrict__ out, int scatter) {
  const ShareLayout L(dm);
  int li = blockIdx.x * blockDim.x + threadIdx.x;
  int count = a.counts[1];
  int bin = -1, gidx = 0;
  if (la.full && la.stage == 0) {
    if (li < a.counts[2]) {
      gidx = a.list1[li];
      const int ne = (int)a.share[(size_t)gidx * a.share_stride + L.hdr];
      bin = lane_bin(ne, ne, 0);
    }
  } else if (la.full) {
    // all worlds but the nominal ones: the front of list0 (position worlds: own row count, the nominal's work) and the back
    const int n0 = a.counts[0];
    if (li < n0 + count) {
      int g0 = li < n0 ? a.list0[li] : a.list0[a.list_cap - 1 - (li - n0)];
      const int nom = a.rowslot[g0 / a.W - a.row_lo];
      const double* shn = a.share + (size_t)nom * a.share_stride;
      const int slot = g0 % a.W;
      if (li < n0) {
        if (slot != 0) { bin = lane_bin((int)a.share[(size_t)li * a.share_stride + L.hdr], (int)shn[L.hdr + 2], 0); gidx = li; }
      } else {
        bin = lane_bin((int)shn[L.hdr], (int)shn[L.hdr + 2], slot <= 2 * a.m);
        gidx = -1 - (li - n0);
      }
    }
  } else if (li < count) {
    gidx = a.list0[a.list_cap - 1 - li];
    const double* sh = a.share + (size_t)(gidx / a.W - a.row_lo) * a.share_stride;
    bin = lane_bin((int)sh[L.hdr], (int)sh[L.hdr + 2], gidx % a.W <= 2 * a.m);
  }
  const int lane = threadIdx.x & 31;
  const unsigned peers = __match_any_sync(0xffffffffu, bin);
  if (bin < 0) return;
  const int leader = __ffs(peers) - 1, rank = __popc(peers & ((1u << lane) - 1u));
  int base = 0;
  if (lane == leader) base = atomicAdd(&la.counts[(scatter ? kLaneBins : 0) + bin], __popc(peers));
  if (!scatter) return;
  base = __shfl_sync(peers, base, leader);
  int off = 0;
  for (int j = 0; j < bin; j++) off += la.counts[j];
  out[off + base + rank] = gidx;
}

#define TOKP_PL(arr, k) (arr)[(k) * 32 + lane]

// sequential line search of one world (MuJoCo PrimalSearch; same evaluation sequence as step_ls.cuh, written the
// way a single thread runs it)
struct LanePnt { double alpha, cost, d0, d1; };

template <int NV>
__device__ __noinline__ LanePnt lane_eval(int ne, const double* __restrict__ jar, const double* __restrict__ jv,
                                             const double* __restrict__ D, int lane, const double qg[3], double alpha,
                                             int& lsiter) {
  lsiter++;
  double q0 = qg[0], q1 = qg[1], q2 = qg[2];
#pragma unroll 2
  for (int r = 0; r < ne; r++) {
    double jr = TOKP_PL(jar, r), v = TOKP_PL(jv, r);
    double x = jr + alpha * v;
    if (x < 0.0) {
      double d = __ldg(&D[r]), DJ0 = d * jr;
      q0 += 0.5 * jr * DJ0;
      q1 += v * DJ0;
      q2 += 0.5 * v * d * v;
    }
  }
  LanePnt p;
  p.alpha = alpha;
  p.cost = alpha * alpha * q2 + alpha * q1 + q0;
  p.d0 = 2.0 * alpha * q2 + q1;
  p.d1 = 2.0 * q2;
  if (p.d1 <= 0.0) p.d1 = tk::kMinVal;
  return p;
}

__device__ __forceinline__ int lane_bracket(LanePnt& p, const LanePnt& c0, const LanePnt& c1, const LanePnt& c2) {
  int flag = 0;
  const LanePnt* c[3] = {&c0, &c1, &c2};
#pragma unroll
  for (int i = 0; i < 3; i++) {
    if (p.d0 < 0.0 && c[i]->d0 < 0.0 && p.d0 < c[i]->d0) { p = *c[i]; flag = 1; }
    else if (p.d0 > 0.0 && c[i]->d0 > 0.0 && p.d0 > c[i]->d0) { p = *c[i]; flag = 2; }
  }
  return flag;
}

// One key-point per warp, one control / velocity world per lane.
template <int NV>
__global__ void __launch_bounds__(32) k_fd_dep_lanes(DM dm, FdArgs a, LaneArgs la) {
  extern __shared__ double smem[];
  constexpr int nv = NV, NTRI = NV * (NV + 1) / 2;
  const int lane = threadIdx.x & 31;
  const int nq = dm.nq, nb = dm.nbody, nu = dm.nu;
  const LanePlan P(dm, la.necap);
  const ShareLayout L(dm);
  double* const vbase = smem + P.vec;
  double* const v_qvel = vbase; double* const v_fsm = vbase + 1 * nv * 32; double* const v_asm = vbase + 2 * nv * 32;
  double* const v_qacc = vbase + 3 * nv * 32; double* const v_Ma = vbase + 4 * nv * 32; double* const v_Mv = vbase + 5 * nv * 32;
  double* const v_grad = vbase + 6 * nv * 32; double* const v_search = vbase + 7 * nv * 32; double* const v_fc = vbase + 8 * nv * 32;
  double* const sc = smem + P.scratch;
  double* const s_cvel = sc; double* const s_cacc = sc + 6 * nb * 32;
  double* const s_cfrc = P.cfrc_in_vec ? v_qacc : sc + 12 * nb * 32;
  double* const s_jar = sc; double* const s_jv = sc + la.necap * 32; double* const s_L = sc + 2 * la.necap * 32;
  int begin = 0, end = 0;
  for (int k = 0; k < kLaneBins; k++) {
    int c = la.counts[k];
    if (k < kLaneHalf) { if (la.heavy) begin += c; else end += c; }
    else if (la.heavy) end += c;
  }
  if (la.heavy) end += begin;
  const bool nominal_stage = la.full && la.stage == 0;
  const int count = end - begin;
  int* cursor = &la.counts[2 * kLaneBins + la.heavy];
  for (;;) {
    int base = 0;
    if (lane == 0) base = atomicAdd(cursor, 32);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= count) break;
    // heaviest bins first (shorter tail)
    const int li = end - 1 - (base + lane);
    const bool active = base + lane < count;
    {
      FdWorld wd;
      wd.kind = 0; wd.col = 0; wd.sgn = 0; wd.slot = 0; wd.row = 0; wd.b = 0; wd.t = 0;
      long long gidx = 0;
      int own = -1;  // full mode: slot of this world's own position stage (nominal / position worlds)
      if (active) {
        if (!la.full) gidx = la.list[li];
        else {
          const int e = la.list[li];
          if (e >= 0) { own = e; gidx = a.list0[e]; }
          else gidx = a.list0[a.list_cap - 1 - (-1 - e)];
        }
        fd_decode(dm, a, gidx, wd);
      }
      const int rowi = (int)(gidx / a.W), row = wd.row, b = wd.b;
      // slot that holds the position stage this world runs on; `shn` = the slot of the key-point's nominal world
      const int nomslot = la.full ? (active ? a.rowslot[rowi - a.row_lo] : 0) : rowi - a.row_lo;
      double* const shn = a.share + (size_t)nomslot * a.share_stride;
      const double* __restrict__ sh = own >= 0 ? a.share + (size_t)own * a.share_stride : shn;
      const bool fullvel = wd.kind != 1;  // control worlds reuse the nominal world's velocity stage
      const double* __restrict__ snap = a.snap + ((size_t)b * (a.T + 1) + wd.t) * a.SD;
      const int ne = active ? (int)__ldg(&sh[L.hdr]) : 0;
      const int hbits = active ? (int)__ldg(&sh[L.hdr + 1]) : 0;
      // the key-point's shared, read-only data (through L1; lanes of the same key-point read the same addresses)
      const double* __restrict__ uM = sh + L.M; const double* __restrict__ uLD = sh + L.LD;
      const double* __restrict__ uLDinv = sh + L.LDinv; const double* __restrict__ uJ = sh + L.J;
      const double* __restrict__ uD = sh + L.D; const double* __restrict__ uaref = shn + L.aref;
      const double* __restrict__ uB = sh + L.Bc; const double* __restrict__ ukterm = sh + L.kterm;
      const double* __restrict__ ufpas = shn + L.fpas; const double* __restrict__ ufbias = shn + L.fbias;
      const double* __restrict__ ucinert = sh + L.cinert; const double* __restrict__ ucdof = sh + L.cdof;
      const double* __restrict__ uLDe = sh + L.LDe; const double* __restrict__ uHinv = sh + L.Hinv;
      const double* __restrict__ uqpos = snap + 1; const double* __restrict__ uqvel = snap + 1 + nq;
      const double* __restrict__ uwarm = snap + 1 + nq + 2 * nv; const double* __restrict__ uctrl = snap + 1 + nq + 3 * nv;
      int st = hbits;
      if (active && ne <= la.necap) {
        const double e = wd.sgn ? -a.eps : a.eps;
        const int pd = wd.kind == 2 ? a.dof_adr[wd.col] : -1;       // perturbed dof
        const int pa = wd.kind == 1 ? a.ctrl_adr[wd.col] : -1;      // perturbed actuator
        // perturbed qpos address (position worlds; hinge / slide joints only on this path: mj_integratePos adds eps)
        const int pqa = wd.kind == 3 ? dm.jnt_qposadr[dm.dof_jntid[a.dof_adr[wd.col]]] : -1;
        int work = 0;
        // ---- velocity stage --------------------------------------------------------------------------------------
#pragma unroll
        for (int i = 0; i < nv; i++) {
          double v = uqvel[i];
          if (i == pd) v += e;
          if (tk::bad_value(v)) st |= tk::ST_BAD_QVEL;
          TOKP_PL(v_qvel, i) = v;
        }
        if (fullvel) {
          // mj_comVel + mj_rne(flg_acc = 0), bodies in tree order (parents first); hinge / slide joints only
          for (int k = 0; k < 6; k++) { TOKP_PL(s_cvel, k) = 0.0; TOKP_PL(s_cfrc, k) = 0.0; }
          TOKP_PL(s_cacc, 0) = 0.0; TOKP_PL(s_cacc, 1) = 0.0; TOKP_PL(s_cacc, 2) = 0.0;
          TOKP_PL(s_cacc, 3) = -dm.gravity[0]; TOKP_PL(s_cacc, 4) = -dm.gravity[1]; TOKP_PL(s_cacc, 5) = -dm.gravity[2];
          for (int i = 1; i < nb; i++) {
            const int p = dm.body_parentid[i];
            double cv[6], tq[6] = {0, 0, 0, 0, 0, 0}, ca[6];
#pragma unroll
            for (int k = 0; k < 6; k++) cv[k] = TOKP_PL(s_cvel, 6 * p + k);
            const int bda = dm.body_dofadr[i], dn = dm.body_dofnum[i];
            for (int j = 0; j < dn; j++) {
              const int dj = bda + j;
              const double* cd = &ucdof[6 * dj];
              const double qv = TOKP_PL(v_qvel, dj);
              double cdd[6];
              tk::cross_motion(cdd, cv, cd);
#pragma unroll
              for (int k = 0; k < 6; k++) { cv[k] += cd[k] * qv; tq[k] += cdd[k] * qv; }
            }
#pragma unroll
            for (int k = 0; k < 6; k++) {
              TOKP_PL(s_cvel, 6 * i + k) = cv[k];
              ca[k] = TOKP_PL(s_cacc, 6 * p + k) + tq[k];
              TOKP_PL(s_cacc, 6 * i + k) = ca[k];
            }
            double t1[6], t2[6], t3[6];
            tk::mul_inert_vec(t1, &ucinert[10 * i], ca);
            tk::mul_inert_vec(t2, &ucinert[10 * i], cv);
            tk::cross_force(t3, cv, t2);
#pragma unroll
            for (int k = 0; k < 6; k++) TOKP_PL(s_cfrc, 6 * i + k) = t1[k] + t3[k];
          }
          for (int i = nb - 1; i > 0; i--) {
            const int p = dm.body_parentid[i];
            if (p)
              for (int k = 0; k < 6; k++) TOKP_PL(s_cfrc, 6 * p + k) += TOKP_PL(s_cfrc, 6 * i + k);
          }
        }
        // qfrc_smooth = passive - bias + actuator
#pragma unroll
        for (int i = 0; i < nv; i++) {
          double fp, fb;
          if (fullvel) {
            fp = -dm.dof_damping[i] * TOKP_PL(v_qvel, i);
            const int bi = dm.dof_bodyid[i];
            const double* cd = &ucdof[6 * i];
            fb = cd[0] * TOKP_PL(s_cfrc, 6 * bi) + cd[1] * TOKP_PL(s_cfrc, 6 * bi + 1) + cd[2] * TOKP_PL(s_cfrc, 6 * bi + 2) +
                 cd[3] * TOKP_PL(s_cfrc, 6 * bi + 3) + cd[4] * TOKP_PL(s_cfrc, 6 * bi + 4) + cd[5] * TOKP_PL(s_cfrc, 6 * bi + 5);
            if (nominal_stage) { shn[L.fpas + i] = fp; shn[L.fbias + i] = fb; }
          } else {
            fp = ufpas[i]; fb = ufbias[i];
          }
          double fa = 0.0;
          for (int ac = 0; ac < nu; ac++) {
            if (dm.actuator_dofadr[ac] != i) continue;
            double ctrl = uctrl[ac];
            if (ac == pa) ctrl += e;
            if (dm.actuator_ctrllimited[ac]) ctrl = fmin(fmax(ctrl, dm.actuator_ctrlrange[2 * ac]), dm.actuator_ctrlrange[2 * ac + 1]);
            double gear = dm.actuator_gear[ac];
            double force = dm.actuator_gainprm[ac] * ctrl;
            if (dm.actuator_biastype[ac] == 1) {
              int qa = dm.jnt_qposadr[dm.actuator_trnid[ac]];
              force += dm.actuator_biasprm[3 * ac] + dm.actuator_biasprm[3 * ac + 1] * (gear * (uqpos[qa] + (qa == pqa ? e : 0.0))) +
                       dm.actuator_biasprm[3 * ac + 2] * (gear * TOKP_PL(v_qvel, i));
            }
            if (dm.actuator_forcelimited[ac]) force = fmin(fmax(force, dm.actuator_forcerange[2 * ac]), dm.actuator_forcerange[2 * ac + 1]);
            fa += gear * force;
          }
          double fs = fp - fb + fa;
          TOKP_PL(v_fsm, i) = fs;
          TOKP_PL(v_asm, i) = fs;
        }
        // qacc_smooth = inv(L'DL) qfrc_smooth (mj_solveLD on the nominal world's factor)
        for (int i = nv - 1; i > 0; i--) {
          double xi = TOKP_PL(v_asm, i);
          if (xi != 0.0)
            for (tk::anc_t<NV> ch = tk::anc_mask<NV>(dm, i) & ~(tk::anc_t<NV>(1) << i); ch;) { const int j = tk::pop_top(ch); TOKP_PL(v_asm, j) -= uLD[i * nv + j] * xi; }
        }
#pragma unroll
        for (int i = 0; i < nv; i++) TOKP_PL(v_asm, i) *= uLDinv[i];
        for (int i = 1; i < nv; i++) {
          double s = TOKP_PL(v_asm, i);
          for (tk::anc_t<NV> ch = tk::anc_mask<NV>(dm, i) & ~(tk::anc_t<NV>(1) << i); ch;) { const int j = tk::pop_top(ch); s -= uLD[i * nv + j] * TOKP_PL(v_asm, j); }
          TOKP_PL(v_asm, i) = s;
        }
        // ---- mj_fwdConstraint: Newton solver ---------------------------------------------------------------------
        if (ne == 0) {
#pragma unroll
          for (int i = 0; i < nv; i++) { TOKP_PL(v_qacc, i) = TOKP_PL(v_asm, i); TOKP_PL(v_fc, i) = 0.0; }
        } else {
          // the velocity-stage scratch is dead from here on: jar / jv / L take its place
          double cw = 0.0, cs = 0.0;
          for (int r = 0; r < ne; r++) {
            const double* Jr = &uJ[r * nv];
            double s1 = 0.0, s2 = 0.0, ar;
            if (fullvel) {
              double vel = 0.0;
#pragma unroll
              for (int i = 0; i < nv; i++) vel += Jr[i] * TOKP_PL(v_qvel, i);
              ar = -uB[r] * vel - ukterm[r];
              if (nominal_stage) shn[L.aref + r] = ar;  // the control worlds of this key-point reuse it
            } else {
              ar = uaref[r];
            }
#pragma unroll
            for (int i = 0; i < nv; i++) { s1 += Jr[i] * uwarm[i]; s2 += Jr[i] * TOKP_PL(v_asm, i); }
            double jw = s1 - ar, js = s2 - ar;
            TOKP_PL(s_jar, r) = jw;
            TOKP_PL(s_jv, r) = js;
            if (jw < 0.0) cw += 0.5 * uD[r] * jw * jw;
            if (js < 0.0) cs += 0.5 * uD[r] * js * js;
          }
          double gs = 0.0;
#pragma unroll
          for (int i = 0; i < nv; i++) {
            double s = 0.0;
#pragma unroll
            for (int j = 0; j < nv; j++) s += uM[i * nv + j] * uwarm[j];
            TOKP_PL(v_Ma, i) = s;
            gs += (s - TOKP_PL(v_fsm, i)) * (uwarm[i] - TOKP_PL(v_asm, i));
          }
          cw += 0.5 * gs;
          const bool use_smooth = cw > cs;
#pragma unroll
          for (int i = 0; i < nv; i++) TOKP_PL(v_qacc, i) = use_smooth ? TOKP_PL(v_asm, i) : uwarm[i];
          if (use_smooth) {
            for (int r = 0; r < ne; r++) TOKP_PL(s_jar, r) = TOKP_PL(s_jv, r);
#pragma unroll
            for (int i = 0; i < nv; i++) {
              double s = 0.0;
#pragma unroll
              for (int j = 0; j < nv; j++) s += uM[i * nv + j] * TOKP_PL(v_qacc, j);
              TOKP_PL(v_Ma, i) = s;
            }
          }
          const double scale = 1.0 / (dm.meaninertia * (double)(nv > 1 ? nv : 1));
          double cost = 0.0, gauss = 0.0, oldcost = 0.0;
          unsigned long long hess_rows = 0ull;
          bool have_factor = false;
          int iter = -1;
          for (;;) {
            // ---- update: forces, cost, qfrc_constraint, gradient, Hessian factor, Newton direction ----
            oldcost = cost;
            double cc = 0.0;
            unsigned long long rows = 0ull;
#pragma unroll
            for (int i = 0; i < nv; i++) TOKP_PL(v_fc, i) = 0.0;
            for (int r = 0; r < ne; r++) {
              double x = TOKP_PL(s_jar, r);
              if (x < 0.0) {
                double d = uD[r];
                cc += 0.5 * d * x * x;
                double f = -d * x;
                if (f != 0.0) {
                  rows |= 1ull << (r & 63);
                  const double* Jr = &uJ[r * nv];
#pragma unroll
                  for (int i = 0; i < nv; i++) TOKP_PL(v_fc, i) += Jr[i] * f;
                }
              }
            }
            double gsum = 0.0;
#pragma unroll
            for (int i = 0; i < nv; i++) {
              double ma = TOKP_PL(v_Ma, i), fs = TOKP_PL(v_fsm, i);
              TOKP_PL(v_grad, i) = ma - fs - TOKP_PL(v_fc, i);
              gsum += (ma - fs) * (TOKP_PL(v_qacc, i) - TOKP_PL(v_asm, i));
            }
            gauss = 0.5 * gsum;
            cost = cc + gauss;
            if (!(have_factor && ne <= 64 && rows == hess_rows)) {
              // H = M + J' diag(D over the rows with a force) J, lower triangle in registers, then Cholesky in place
              double h[NTRI];
#pragma unroll
              for (int i = 0; i < nv; i++)
#pragma unroll
                for (int j = 0; j <= i; j++) h[i * (i + 1) / 2 + j] = uM[i * nv + j];
              for (int r = 0; r < ne; r++) {
                double x = TOKP_PL(s_jar, r);
                if (!(x < 0.0) || -uD[r] * x == 0.0) continue;
                const double* Jr = &uJ[r * nv];
                const double d = uD[r];
#pragma unroll
                for (int i = 0; i < nv; i++) {
                  const double dji = d * Jr[i];
#pragma unroll
                  for (int j = 0; j <= i; j++) h[i * (i + 1) / 2 + j] += dji * Jr[j];
                }
              }
#pragma unroll
              for (int j = 0; j < nv; j++) {
                double dj = h[j * (j + 1) / 2 + j];
                if (dj < tk::kMinVal) dj = tk::kMinVal;
                dj = sqrt(dj);
                const double inv = 1.0 / dj;
                h[j * (j + 1) / 2 + j] = dj;
#pragma unroll
                for (int i = j + 1; i < nv; i++) h[i * (i + 1) / 2 + j] *= inv;
#pragma unroll
                for (int i = j + 1; i < nv; i++)
#pragma unroll
                  for (int k = j + 1; k <= i; k++) h[i * (i + 1) / 2 + k] -= h[i * (i + 1) / 2 + j] * h[k * (k + 1) / 2 + j];
              }
#pragma unroll
              for (int k = 0; k < NTRI; k++) TOKP_PL(s_L, k) = h[k];
            }
            hess_rows = rows;
            have_factor = true;
            // Mgrad = inv(L L') grad; search = -Mgrad
            for (int i = 0; i < nv; i++) {
              double tt = TOKP_PL(v_grad, i);
              for (int k = 0; k < i; k++) tt -= TOKP_PL(s_L, i * (i + 1) / 2 + k) * TOKP_PL(v_search, k);
              TOKP_PL(v_search, i) = tt / TOKP_PL(s_L, i * (i + 1) / 2 + i);
            }
            for (int i = nv - 1; i >= 0; i--) {
              double tt = TOKP_PL(v_search, i);
              for (int k = i + 1; k < nv; k++) tt -= TOKP_PL(s_L, k * (k + 1) / 2 + i) * TOKP_PL(v_search, k);
              TOKP_PL(v_search, i) = tt / TOKP_PL(s_L, i * (i + 1) / 2 + i);
            }
            double gn = 0.0, sn = 0.0;
#pragma unroll
            for (int i = 0; i < nv; i++) {
              double gi = TOKP_PL(v_grad, i), si = -TOKP_PL(v_search, i);
              TOKP_PL(v_search, i) = si;
              gn += gi * gi;
              sn += si * si;
            }
            if (iter >= 0) {
              double improvement = scale * (oldcost - cost), gradient = scale * sqrt(gn);
              if (improvement < a.tolerance || gradient < a.tolerance) break;
            }
            iter++;
            if (iter >= a.iterations) break;
            // ---- exact line search along `search` ----
            double alpha = 0.0;
            const double snorm = sqrt(sn);
            enum { S_P0 = 0, S_P1, S_ONESIDED, S_P1NEXT, S_MID, S_UB1, S_UB2, S_DONE };
            int state = S_DONE;
            double gtol = 0.0, qg0 = 0.0, qg1 = 0.0, qg2 = 0.0;
            if (!(snorm < tk::kMinVal)) {
              state = S_P0;
              gtol = a.tolerance * dm.ls_tolerance * snorm / scale;
              double qa = 0.0, qb = 0.0, qc = 0.0;
#pragma unroll
              for (int i = 0; i < nv; i++) {
                double s = 0.0;
#pragma unroll
                for (int j = 0; j < nv; j++) s += uM[i * nv + j] * TOKP_PL(v_search, j);
                TOKP_PL(v_Mv, i) = s;
                double si = TOKP_PL(v_search, i);
                qa += si * TOKP_PL(v_Ma, i);
                qb += TOKP_PL(v_fsm, i) * si;
                qc += si * s;
              }
              for (int r = 0; r < ne; r++) {
                const double* Jr = &uJ[r * nv];
                double s = 0.0;
#pragma unroll
                for (int i = 0; i < nv; i++) s += Jr[i] * TOKP_PL(v_search, i);
                TOKP_PL(s_jv, r) = s;
              }
              qg0 = gauss; qg1 = qa - qb; qg2 = 0.5 * qc;
            }
            {
              const int LS = dm.ls_iterations;
              // PrimalSearch as a state machine with ONE cost evaluation per trip (same evaluation sequence as the
              // sequential algorithm, see step_ls.cuh): lanes in different phases of their searches still evaluate
              // together, so the row loop below - the bulk of the search - runs with all lanes of the warp.  The vote
              // in the loop condition keeps the lanes that entered together in the same trip (and the compiler from
              // specialising the first trips into separate copies of the evaluation).
              int lsiter = 0, p2update = 0, b2 = 0;
              double dir = 0.0, a_eval = 0.0;
              LanePnt p0 = {0, 0, 0, 0}, p1 = p0, p2 = p0, p1next = p0, p2next = p0;
              const unsigned lsmask = __activemask();
              while (__any_sync(lsmask, state != S_DONE)) {
                if (state == S_DONE) continue;
                LanePnt pe;
                {
                  lsiter++;
                  double q0 = qg0, q1 = qg1, q2 = qg2;
#pragma unroll 2
                  for (int r = 0; r < ne; r++) {
                    double jr = TOKP_PL(s_jar, r), v = TOKP_PL(s_jv, r);
                    double x = jr + a_eval * v;
                    if (x < 0.0) {
                      double d = uD[r], DJ0 = d * jr;
                      q0 += 0.5 * jr * DJ0;
                      q1 += v * DJ0;
                      q2 += 0.5 * v * d * v;
                    }
                  }
                  pe.alpha = a_eval;
                  pe.cost = a_eval * a_eval * q2 + a_eval * q1 + q0;
                  pe.d0 = 2.0 * a_eval * q2 + q1;
                  pe.d1 = 2.0 * q2;
                  if (pe.d1 <= 0.0) pe.d1 = tk::kMinVal;
                }
                enum { C_NONE = 0, C_ONESIDED, C_AFTER, C_BRACKET, C_FINAL };
                enum { N_NONE = 0, N_P0, N_P1, N_P2 };
                int ctl = C_NONE, nsel = N_NONE;
                bool fixed_point_check = false;
                if (state == S_P0) {
                  p0 = pe; nsel = N_P0; state = S_P1;
                } else if (state == S_P1) {
                  p1 = pe;
                  if (p0.cost < p1.cost) p1 = p0;
                  if (fabs(p1.d0) < gtol) { alpha = p1.alpha; state = S_DONE; }
                  else { dir = (p1.d0 < 0.0) ? 1.0 : -1.0; p2update = 0; p2 = p1; ctl = C_ONESIDED; }
                } else if (state == S_ONESIDED) {
                  p1 = pe;
                  if (fabs(p1.d0) < gtol) { alpha = p1.alpha; state = S_DONE; }
                  else ctl = C_ONESIDED;
                } else if (state == S_P1NEXT) {
                  p1next = pe; ctl = C_BRACKET;
                } else if (state == S_MID) {
                  const LanePnt c0 = p1next, c1 = p2next, c2 = pe;
                  double bestcost = 0.0, bestalpha = 0.0;
                  bool found = false;
                  if (fabs(c0.d0) < gtol) { bestcost = c0.cost; bestalpha = c0.alpha; found = true; }
                  if (fabs(c1.d0) < gtol && (!found || c1.cost < bestcost)) { bestcost = c1.cost; bestalpha = c1.alpha; found = true; }
                  if (fabs(c2.d0) < gtol && (!found || c2.cost < bestcost)) { bestcost = c2.cost; bestalpha = c2.alpha; found = true; }
                  if (found) { alpha = bestalpha; state = S_DONE; }
                  else {
                    int b1 = lane_bracket(p1, c0, c1, c2);
                    b2 = lane_bracket(p2, c0, c1, c2);
                    if (b1) { nsel = N_P1; state = S_UB1; }
                    else if (b2) { nsel = N_P2; state = S_UB2; }
                    else { alpha = pe.alpha; state = S_DONE; }
                  }
                } else if (state == S_UB1) {
                  p1next = pe;
                  if (b2) { nsel = N_P2; state = S_UB2; }
                  else ctl = C_BRACKET;
                } else {  // S_UB2
                  p2next = pe; ctl = C_BRACKET;
                }
                if (ctl == C_ONESIDED) {
                  if (p1.d0 * dir <= -gtol && lsiter < LS) {
                    p2 = p1; p2update = 1; nsel = N_P1; fixed_point_check = true; state = S_ONESIDED;
                  } else ctl = C_AFTER;
                }
                if (ctl == C_AFTER) {
                  if (lsiter >= LS || !p2update) { alpha = p1.alpha; state = S_DONE; }
                  else { p2next = p1; nsel = N_P1; state = S_P1NEXT; }
                }
                if (ctl == C_BRACKET) {
                  if (lsiter < LS) { a_eval = 0.5 * (p1.alpha + p2.alpha); state = S_MID; }
                  else ctl = C_FINAL;
                }
                if (ctl == C_FINAL) {
                  if (p1.cost <= p2.cost && p1.cost < p0.cost) alpha = p1.alpha;
                  else if (p2.cost <= p1.cost && p2.cost < p0.cost) alpha = p2.alpha;
                  else alpha = 0.0;
                  state = S_DONE;
                }
                if (nsel != N_NONE) {
                  const LanePnt& ps = nsel == N_P0 ? p0 : (nsel == N_P1 ? p1 : p2);
                  a_eval = ps.alpha - ps.d0 / ps.d1;
                  // floating-point fixed point of the one-sided Newton iteration: every further evaluation would return
                  // p1 again until the iteration cap and p1.alpha would be returned (see step_ls.cuh)
                  if (fixed_point_check && a_eval == p1.alpha) { alpha = p1.alpha; state = S_DONE; }
                }
              }
              work += lsiter;
            }
            if (alpha == 0.0) break;
#pragma unroll
            for (int i = 0; i < nv; i++) {
              TOKP_PL(v_qacc, i) += alpha * TOKP_PL(v_search, i);
              TOKP_PL(v_Ma, i) += alpha * TOKP_PL(v_Mv, i);
            }
            for (int r = 0; r < ne; r++) TOKP_PL(s_jar, r) += alpha * TOKP_PL(s_jv, r);
          }
        }
        if (nominal_stage) shn[L.hdr + 2] = (double)work;  // sort key of the other worlds of this key-point
        // ---- mj_EulerSkip(skipfactor) + mj_advance ---------------------------------------------------------------
#pragma unroll
        for (int i = 0; i < nv; i++) if (tk::bad_value(TOKP_PL(v_qacc, i))) st |= tk::ST_BAD_QACC;
        if (dm.any_damping) {
#pragma unroll
          for (int i = 0; i < nv; i++) TOKP_PL(v_grad, i) = TOKP_PL(v_fsm, i) + TOKP_PL(v_fc, i);
          for (int i = nv - 1; i > 0; i--) {
            double xi = TOKP_PL(v_grad, i);
            if (xi != 0.0)
              for (tk::anc_t<NV> ch = tk::anc_mask<NV>(dm, i) & ~(tk::anc_t<NV>(1) << i); ch;) { const int j = tk::pop_top(ch); TOKP_PL(v_grad, j) -= uLDe[i * nv + j] * xi; }
          }
#pragma unroll
          for (int i = 0; i < nv; i++) TOKP_PL(v_grad, i) *= uHinv[i];
          for (int i = 1; i < nv; i++) {
            double s = TOKP_PL(v_grad, i);
            for (tk::anc_t<NV> ch = tk::anc_mask<NV>(dm, i) & ~(tk::anc_t<NV>(1) << i); ch;) { const int j = tk::pop_top(ch); s -= uLDe[i * nv + j] * TOKP_PL(v_grad, j); }
            TOKP_PL(v_grad, i) = s;
          }
        } else {
#pragma unroll
          for (int i = 0; i < nv; i++) TOKP_PL(v_grad, i) = TOKP_PL(v_qacc, i);
        }
        const double hstep = dm.timestep;
        double* o = a.fd_out + ((size_t)row * a.W + wd.slot) * (nq + nv);
#pragma unroll
        for (int i = 0; i < nv; i++) {
          double v = TOKP_PL(v_qvel, i) + TOKP_PL(v_grad, i) * hstep;
          TOKP_PL(v_qvel, i) = v;
          o[nq + i] = v;
        }
        for (int j = 0; j < dm.njnt; j++) {  // hinge / slide joints only (checked when the lane path is selected)
          int pq = dm.jnt_qposadr[j], va = dm.jnt_dofadr[j];
          double q = uqpos[pq] + (pq == pqa ? e : 0.0);
          if (tk::bad_value(q)) st |= tk::ST_BAD_QPOS;
          o[pq] = q + hstep * TOKP_PL(v_qvel, va);
        }
      } else if (active) {
        st |= tk::ST_OVERFLOW;  // more constraint rows than the lane workspace holds
      }
      if (active) {
        a.status[(size_t)row * a.W + wd.slot] = (st & 63) | (1 << 30);
        if (st & 63) atomicOr(&a.wstatus[2 * b], st & 63);
      }
    }
    __syncwarp();
  }
}

#undef TOKP_PL

}  // namespace tokp
