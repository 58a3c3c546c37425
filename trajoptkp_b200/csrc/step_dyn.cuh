// Velocity / acceleration stages, Newton constraint solver with exact line search and the
// semi-implicit Euler update of the batched step (MuJoCo: mj_fwdVelocity, mj_fwdActuation,
// mj_fwdAcceleration, mj_fwdConstraint/mj_solNewton, mj_EulerSkip).
#pragma once
#include "step_con.cuh"

namespace tk {

// mj_comVel and the forward pass of mj_rne(flg_acc = 0) in ONE level-synchronous walk down the tree: a body's cvel and
// cdof_dot, then its cacc = cacc[parent] + cdof_dot qvel and cfrc = I cacc + cvel x* (I cvel), all of which need only the
// parent's values of the previous level.  Same operations per entry as two separate walks, one barrier per level less.
template <int G, int NV>
TK_DEV void com_vel(const DM& m, WS& w, const Grp<G, NV>& g) {
  if (g.gl == 0) {
    for (int k = 0; k < 6; k++) w.cvel[k] = 0.0;
    w.cacc[0] = w.cacc[1] = w.cacc[2] = 0.0;
    w.cacc[3] = -m.gravity[0]; w.cacc[4] = -m.gravity[1]; w.cacc[5] = -m.gravity[2];
    for (int k = 0; k < 6; k++) w.cfrc[k] = 0.0;
  }
  g.sync();
  for (int depth = 1; depth <= m.treedepth; depth++) {
    for (int i = 1 + g.gl; i < m.nbody; i += G) {
      if (m.body_depth[i] != depth) continue;
      const int pid = m.body_parentid[i];
      double cv[6];
      for (int k = 0; k < 6; k++) cv[k] = w.cvel[6 * pid + k];
      int bda = m.body_dofadr[i], dn = m.body_dofnum[i];
      int j = 0;
      while (j < dn) {
        int dj = bda + j;
        if (m.jnt_type[m.dof_jntid[dj]] == JNT_FREE) {
          for (int k = 0; k < 18; k++) w.cdof_dot[6 * dj + k] = 0.0;
          for (int k = 0; k < 3; k++)
            for (int c = 0; c < 6; c++) cv[c] += w.cdof[6 * (dj + k) + c] * w.qvel[dj + k];
          dj += 3;
          for (int k = 0; k < 3; k++) cross_motion(&w.cdof_dot[6 * (dj + k)], cv, &w.cdof[6 * (dj + k)]);
          for (int k = 0; k < 3; k++)
            for (int c = 0; c < 6; c++) cv[c] += w.cdof[6 * (dj + k) + c] * w.qvel[dj + k];
          j += 6;
        } else {
          cross_motion(&w.cdof_dot[6 * dj], cv, &w.cdof[6 * dj]);
          for (int c = 0; c < 6; c++) cv[c] += w.cdof[6 * dj + c] * w.qvel[dj];
          j += 1;
        }
      }
      for (int k = 0; k < 6; k++) w.cvel[6 * i + k] = cv[k];
      // mj_rne forward pass for this body
      double t[6] = {0, 0, 0, 0, 0, 0};
      for (int jj = 0; jj < dn; jj++)
        for (int c = 0; c < 6; c++) t[c] += w.cdof_dot[6 * (bda + jj) + c] * w.qvel[bda + jj];
      double ca[6];
      for (int c = 0; c < 6; c++) { ca[c] = w.cacc[6 * pid + c] + t[c]; w.cacc[6 * i + c] = ca[c]; }
      double t1[6], t2[6], t3[6];
      mul_inert_vec(t1, &w.cinert[10 * i], ca);
      mul_inert_vec(t2, &w.cinert[10 * i], cv);
      cross_force(t3, cv, t2);
      for (int c = 0; c < 6; c++) w.cfrc[6 * i + c] = t1[c] + t3[c];
    }
    g.sync();
  }
}

// mj_passive (damping + gravity compensation) and the backward pass of mj_rne (cacc / cfrc come from com_vel's walk)
template <int G, int NV>
TK_DEV void passive_and_bias(const DM& m, WS& w, const Grp<G, NV>& g) {
  const int nv = NV ? NV : m.nv;
  for (int i = g.gl; i < nv; i += G) {
    double f = -m.dof_damping[i] * w.qvel[i];
    if (m.any_gravcomp) {
      for (int b = 1; b < m.nbody; b++) {
        double gc = m.body_gravcomp[b];
        if (gc == 0.0) continue;
        double col[3];
        jac_col(m, w, col, &w.xipos[3 * b], b, i);
        double s = m.body_mass[b] * gc;
        f += col[0] * (-m.gravity[0] * s) + col[1] * (-m.gravity[1] * s) + col[2] * (-m.gravity[2] * s);
      }
    }
    w.frc_passive[i] = f;
  }
  // bias[i] = cdof_i . (sum of body forces over the subtree of the dof's body)
  for (int i = g.gl; i < nv; i += G) {
    int b = m.dof_bodyid[i];
    unsigned lo = m.body_subtreemask_lo[b], hi = m.body_subtreemask_hi[b];
    double f[6] = {0, 0, 0, 0, 0, 0};
    for (int j = m.nbody - 1; j >= b; j--) {
      if (!in_mask(lo, hi, j)) continue;
      for (int c = 0; c < 6; c++) f[c] += w.cfrc[6 * j + c];
    }
    const double* cd = &w.cdof[6 * i];
    w.frc_bias[i] = cd[0] * f[0] + cd[1] * f[1] + cd[2] * f[2] + cd[3] * f[3] + cd[4] * f[4] + cd[5] * f[5];
  }
  g.sync();
}

// mj_fwdActuation + mj_fwdAcceleration
template <int G, int NV>
TK_DEV void actuation_acceleration(const DM& m, WS& w, const Grp<G, NV>& g) {
  const int nv = NV ? NV : m.nv;
  for (int i = g.gl; i < nv; i += G) {
    double fa = 0.0;
    for (int a = 0; a < m.nu; a++) {
      if (m.actuator_dofadr[a] != i) continue;
      double ctrl = w.ctrl[a];
      if (m.actuator_ctrllimited[a]) ctrl = fmin(fmax(ctrl, m.actuator_ctrlrange[2 * a]), m.actuator_ctrlrange[2 * a + 1]);
      double gear = m.actuator_gear[a];
      double force = m.actuator_gainprm[a] * ctrl;
      if (m.actuator_biastype[a] == 1) {
        int qa = m.jnt_qposadr[m.actuator_trnid[a]];
        force += m.actuator_biasprm[3 * a] + m.actuator_biasprm[3 * a + 1] * (gear * w.qpos[qa]) +
                 m.actuator_biasprm[3 * a + 2] * (gear * w.qvel[i]);
      }
      if (m.actuator_forcelimited[a]) force = fmin(fmax(force, m.actuator_forcerange[2 * a]), m.actuator_forcerange[2 * a + 1]);
      fa += gear * force;
    }
    w.frc_act[i] = fa;
    double fs = w.frc_passive[i] - w.frc_bias[i] + fa;
    w.frc_smooth[i] = fs;
    w.acc_smooth[i] = fs;
  }
  g.sync();
  solve_ldl(m, w.LD, w.LDinv, w.acc_smooth, g);
}

// ---- Newton solver --------------------------------------------------------------------------------
struct Pnt { double alpha, cost, d0, d1; };

// cost of the active rows at jar (s(jar)); optionally writes forces
template <int G, int NV>
TK_DEV double constraint_cost(const WS& w, const double* jar, const Grp<G, NV>& g) {
  double c = 0.0;
  for (int r = g.gl; r < w.nefc; r += G) {
    double x = jar[r];
    if (x < 0.0) c += 0.5 * w.efc_D[r] * x * x;
  }
  return g.sum(c);
}

template <int G, int NV>
TK_DEV Pnt primal_eval(const WS& w, const Grp<G, NV>& g, const double qg[3], double alpha, int& lsiter) {
  lsiter++;
  double q0 = 0.0, q1 = 0.0, q2 = 0.0;
  for (int r = g.gl; r < w.nefc; r += G) {
    double x = w.jar[r] + alpha * w.jv[r];
    if (x < 0.0) { q0 += w.quad[3 * r]; q1 += w.quad[3 * r + 1]; q2 += w.quad[3 * r + 2]; }
  }
  q0 = g.sum(q0) + qg[0]; q1 = g.sum(q1) + qg[1]; q2 = g.sum(q2) + qg[2];
  Pnt p;
  p.alpha = alpha;
  p.cost = alpha * alpha * q2 + alpha * q1 + q0;
  p.d0 = 2.0 * alpha * q2 + q1;
  p.d1 = 2.0 * q2;
  if (p.d1 <= 0.0) p.d1 = kMinVal;
  return p;
}

// Lower-triangle entries of an nv x nv matrix owned by one lane (entry idx = lane + k*G, row-major over the triangle):
// used by the Hessian assembly and the Cholesky trailing update when nv(nv+1)/2 <= TRI_MAX*G.
__host__ __device__ constexpr int tri_max(int lanes, int nv) { return nv ? (nv * (nv + 1) / 2 + lanes - 1) / lanes : 8; }
template <int G, int NV>
struct TriMap {
  static constexpr int TRI_MAX = tri_max(G, NV);
  int cnt;             // entries of this lane
  int ei[TRI_MAX], ej[TRI_MAX];
  TK_DEV void init(int nv, int gl) {
    int ntri = nv * (nv + 1) / 2;
    cnt = 0;
#pragma unroll
    for (int k = 0; k < TRI_MAX; k++) {
      int idx = gl + k * G;
      ei[k] = 0; ej[k] = 0;
      if (idx < ntri) {
        int i = (int)((sqrt(8.0 * (double)idx + 1.0) - 1.0) * 0.5);
        while (i * (i + 1) / 2 > idx) i--;
        while ((i + 1) * (i + 2) / 2 <= idx) i++;
        ei[k] = i; ej[k] = idx - i * (i + 1) / 2;
        cnt = k + 1;
      }
    }
  }
};

// same factorisation as chol_factor with the trailing update driven by the lane's triangle entries (no integer
// division, no wasted upper-triangle iterations); identical arithmetic per entry
template <int G, int NV>
TK_DEV void chol_factor_tri(double* H, int n, const Grp<G, NV>& g, const TriMap<G, NV>& tm) {
  constexpr int TRI_MAX = TriMap<G, NV>::TRI_MAX;
#pragma unroll 1
  for (int j = 0; j < n; j++) {
    double d = H[j * n + j];
    if (d < kMinVal) d = kMinVal;
    d = sqrt(d);
    double inv = 1.0 / d;
    g.sync();
    for (int i = j + g.gl; i < n; i += G) H[i * n + j] = (i == j) ? d : H[i * n + j] * inv;
    g.sync();
#pragma unroll
    for (int k = 0; k < TRI_MAX; k++)
      if (k < tm.cnt && tm.ej[k] > j) H[tm.ei[k] * n + tm.ej[k]] -= H[tm.ei[k] * n + j] * H[tm.ej[k] * n + j];
    g.sync();
  }
}

// Walk over the entries of a lower triangle in row-major order (entry e = row * (row + 1) / 2 + col, col <= row), G
// entries at a time: no integer division, no iterations spent on the upper triangle.
struct TriWalk {
  int row, col;
  TK_DEV explicit TriWalk(int first) : row(0), col(first) { settle(); }
  TK_DEV void settle() { while (col > row) { col -= row + 1; row++; } }
  TK_DEV void advance(int by) { col += by; settle(); }
};

// in-place dense Cholesky (lower) of the nv x nv matrix in w.H, right-looking, lanes over the rows of the trailing block
template <int G, int NV>
TK_DEV void chol_factor(double* H, int n, const Grp<G, NV>& g) {
#pragma unroll 1
  for (int j = 0; j < n; j++) {
    double d = H[j * n + j];
    if (d < kMinVal) d = kMinVal;
    d = sqrt(d);
    double inv = 1.0 / d;
    g.sync();
    for (int i = j + g.gl; i < n; i += G) H[i * n + j] = (i == j) ? d : H[i * n + j] * inv;
    g.sync();
    // trailing update, a lane per row i of the block below / right of column j: H[i][k] -= L[i][j] L[k][j] for
    // j < k <= i.  No index arithmetic; column j is not written here, so four entries are loaded ahead of their stores
    for (int i = j + 1 + g.gl; i < n; i += G) {
      const double lij = H[i * n + j];
      double* Hi = &H[i * n];
      int k = j + 1;
      for (; k + 3 <= i; k += 4) {
        const double a0 = Hi[k], a1 = Hi[k + 1], a2 = Hi[k + 2], a3 = Hi[k + 3];
        const double b0 = H[k * n + j], b1 = H[(k + 1) * n + j], b2 = H[(k + 2) * n + j], b3 = H[(k + 3) * n + j];
        Hi[k] = a0 - lij * b0; Hi[k + 1] = a1 - lij * b1; Hi[k + 2] = a2 - lij * b2; Hi[k + 3] = a3 - lij * b3;
      }
      for (; k <= i; k++) Hi[k] -= lij * H[k * n + j];
    }
    g.sync();
  }
}

// x = inv(L L') b  (column-oriented substitutions, lanes over the remaining rows)
template <int G, int NV>
TK_DEV void chol_solve(const double* L, int n, double* x, const double* b, const Grp<G, NV>& g) {
  for (int i = g.gl; i < n; i += G) x[i] = b[i];
  g.sync();
#pragma unroll 1
  for (int j = 0; j < n; j++) {
    double xj = x[j] / L[j * n + j];
    g.sync();
    if (g.gl == 0) x[j] = xj;
    for (int i = j + 1 + g.gl; i < n; i += G) x[i] -= L[i * n + j] * xj;
    g.sync();
  }
#pragma unroll 1
  for (int j = n - 1; j >= 0; j--) {
    double xj = x[j] / L[j * n + j];
    g.sync();
    if (g.gl == 0) x[j] = xj;
    for (int i = g.gl; i < j; i += G) x[i] -= L[j * n + i] * xj;
    g.sync();
  }
}

// The same solve for n <= G with the vector in registers (lane i holds x[i]) and the pivots broadcast by shuffles: no
// barrier and no shared-memory round trip per column, the reciprocal pivots are formed once, off the dependent chain.
// Same elimination order as chol_solve; the result differs from it only by the rounding of x * (1 / l) against x / l.
template <int G, int NV>
TK_DEV void chol_solve_reg(const double* L, int n, double* x, const double* b, const Grp<G, NV>& g) {
  const int i = g.gl;
  const bool mine = i < n;
  const int ic = mine ? i : 0;  // a valid row / column for the lanes past n (their loads are not used)
  double xi = mine ? b[i] : 0.0;
  const double dinv = mine ? 1.0 / L[i * n + i] : 0.0;
  // the entry of L for the next trip is loaded before this trip's shuffle: it is off the dependent chain
  double lnext = L[ic * n];
#pragma unroll 1
  for (int j = 0; j < n; j++) {
    const double l = lnext;
    if (j + 1 < n) lnext = L[ic * n + j + 1];
    const double xj = __shfl_sync(g.mask, xi * dinv, j, G);
    if (i == j) xi = xj;
    else if (mine && i > j) xi -= l * xj;
  }
  lnext = L[(n - 1) * n + ic];
#pragma unroll 1
  for (int j = n - 1; j >= 0; j--) {
    const double l = lnext;
    if (j > 0) lnext = L[(j - 1) * n + ic];
    const double xj = __shfl_sync(g.mask, xi * dinv, j, G);
    if (i == j) xi = xj;
    else if (i < j) xi -= l * xj;
  }
  if (mine) x[i] = xi;
  g.sync();
}

}  // namespace tk
#include "step_ls.cuh"
namespace tk {

// mj_fwdConstraint with the Newton solver.  The Newton and line-search loops are warp-uniform (Grp::any_warp): every
// running world of the warp passes through them together; worlds without constraints or already converged idle.
// chol_solve_reg for G < n <= 2 G: two entries of x per lane (rows gl and gl + G)
template <int G, int NV>
TK_DEV void chol_solve_reg2(const double* L, int n, double* x, const double* b, const Grp<G, NV>& g) {
  const int i0 = g.gl, i1 = g.gl + G;
  const bool m1 = i1 < n;  // i0 < G < n always
  double x0 = b[i0], x1 = m1 ? b[i1] : 0.0;
  const double d0 = 1.0 / L[i0 * n + i0], d1 = m1 ? 1.0 / L[i1 * n + i1] : 0.0;
  const int i1c = m1 ? i1 : i0;
  double l0n = L[i0 * n], l1n = L[i1c * n];  // next trip's entries, loaded ahead of the shuffle
#pragma unroll 1
  for (int j = 0; j < n; j++) {
    const double l0 = l0n, l1 = l1n;
    if (j + 1 < n) { l0n = L[i0 * n + j + 1]; l1n = L[i1c * n + j + 1]; }
    const bool lowj = j < G;
    const double xj = __shfl_sync(g.mask, lowj ? x0 * d0 : x1 * d1, j & (G - 1), G);
    if (lowj) { if (i0 == j) x0 = xj; else if (i0 > j) x0 -= l0 * xj; }
    if (!lowj && i1 == j) x1 = xj;
    else if (m1 && i1 > j) x1 -= l1 * xj;
  }
  l0n = L[(n - 1) * n + i0]; l1n = L[(n - 1) * n + i1c];
#pragma unroll 1
  for (int j = n - 1; j >= 0; j--) {
    const double l0 = l0n, l1 = l1n;
    if (j > 0) { l0n = L[(j - 1) * n + i0]; l1n = L[(j - 1) * n + i1c]; }
    const bool lowj = j < G;
    const double xj = __shfl_sync(g.mask, lowj ? x0 * d0 : x1 * d1, j & (G - 1), G);
    if (!lowj) { if (i1 == j) x1 = xj; else if (i1 < j) x1 -= l1 * xj; }
    if (lowj && i0 == j) x0 = xj;
    else if (i0 < j) x0 -= l0 * xj;
  }
  x[i0] = x0;
  if (m1) x[i1] = x1;
  g.sync();
}

template <int G, int NV>
TK_DEV void fwd_constraint(const DM& m, WS& w, const Grp<G, NV>& g, int iterations, double tolerance) {
  const int nv = NV ? NV : m.nv, nefc = w.nefc;
  const bool has = nefc > 0;
  double cost = 0.0, gauss = 0.0;
  constexpr int TRI_MAX = TriMap<G, NV>::TRI_MAX;
  const bool tri = nv * (nv + 1) / 2 <= TRI_MAX * G;
  TriMap<G, NV> tm;
  tm.cnt = 0;
  if (tri && has) tm.init(nv, g.gl);
  // rows of this lane that entered the last assembled Hessian (bit k = row gl + k*G); 64 bits cover maxefc <= 64*G
  unsigned long long hess_rows = 0ull;
  bool have_factor = false;
  auto update = [&]() {
    // forces, cost, qfrc_constraint, gradient, Hessian, Newton direction
    double c = 0.0;
    unsigned long long rows = 0ull;
    int kbit = 0;
    for (int r = g.gl; r < nefc; r += G, kbit++) {
      double x = w.jar[r];
      if (x < 0.0) { c += 0.5 * w.efc_D[r] * x * x; w.efc_force[r] = -w.efc_D[r] * x; }
      else w.efc_force[r] = 0.0;
      if (w.efc_force[r] != 0.0) rows |= 1ull << (kbit & 63);
    }
    // H = M + J' diag(D over the rows with a force) J depends on the state only through that row set: when the set is
    // the one of the previous Newton iteration the factor in w.H is still the factor of this H, bit for bit
    const bool same_factor = have_factor && nefc <= 64 * G && g.all(rows == hess_rows);
    hess_rows = rows;
    have_factor = true;
    c = g.sum(c);
    g.sync();
    double gsum = 0.0;
    for (int i = g.gl; i < nv; i += G) {
      double fc = 0.0;
      for (int r = 0; r < nefc; r++) fc += w.J[r * nv + i] * w.efc_force[r];
      w.frc_con[i] = fc;
      w.grad[i] = w.Ma[i] - w.frc_smooth[i] - fc;
      gsum += (w.Ma[i] - w.frc_smooth[i]) * (w.qacc[i] - w.acc_smooth[i]);
    }
    gauss = 0.5 * g.sum(gsum);
    cost = c + gauss;
    TK_CLK(w, g, 7); TK_CNT(w, g, 20);
    if (same_factor) {
      // keep the factor
    } else if (tri) {
      // rows outermost, the lane's triangle entries in registers: inactive rows cost one broadcast load
      double h[TRI_MAX];
#pragma unroll
      for (int k = 0; k < TRI_MAX; k++) h[k] = k < tm.cnt ? w.M[tm.ei[k] * nv + tm.ej[k]] : 0.0;
      for (int r = 0; r < nefc; r++) {
        if (w.efc_force[r] == 0.0) continue;
        double D = w.efc_D[r];
        const double* Jr = &w.J[r * nv];
#pragma unroll
        for (int k = 0; k < TRI_MAX; k++)
          if (k < tm.cnt) h[k] += D * Jr[tm.ei[k]] * Jr[tm.ej[k]];
      }
#pragma unroll
      for (int k = 0; k < TRI_MAX; k++)
        if (k < tm.cnt) w.H[tm.ei[k] * nv + tm.ej[k]] = h[k];
      g.sync();
      TK_CLK(w, g, 8); TK_CNT(w, g, 21);
      chol_factor_tri(w.H, nv, g, tm);
    } else {
      // four entries of the triangle per lane and pass over the rows (independent accumulators)
      const int ntri = nv * (nv + 1) / 2;
      TriWalk t(g.gl);
      for (int e = g.gl; e < ntri; e += 4 * G) {
        int ei[4], ej[4];
        double h[4];
#pragma unroll
        for (int b = 0; b < 4; b++) {
          const bool ok = e + b * G < ntri;
          ei[b] = ok ? t.row : -1; ej[b] = ok ? t.col : 0;
          h[b] = ok ? w.M[t.row * nv + t.col] : 0.0;
          t.advance(G);
        }
        for (int r = 0; r < nefc; r++) {
          if (w.efc_force[r] == 0.0) continue;
          const double D = w.efc_D[r];
          const double* Jr = &w.J[r * nv];
#pragma unroll
          for (int b = 0; b < 4; b++) h[b] += D * Jr[ei[b] < 0 ? 0 : ei[b]] * Jr[ej[b]];
        }
#pragma unroll
        for (int b = 0; b < 4; b++)
          if (ei[b] >= 0) w.H[ei[b] * nv + ej[b]] = h[b];
      }
      g.sync();
      TK_CLK(w, g, 8); TK_CNT(w, g, 21);
      chol_factor(w.H, nv, g);
    }
    TK_CLK(w, g, 9);
    if (nv <= G) chol_solve_reg(w.H, nv, w.Mgrad, w.grad, g);
    else if (nv <= 2 * G) chol_solve_reg2(w.H, nv, w.Mgrad, w.grad, g);
    else chol_solve(w.H, nv, w.Mgrad, w.grad, g);
    for (int i = g.gl; i < nv; i += G) w.search[i] = -w.Mgrad[i];
    g.sync();
    TK_CLK(w, g, 10);
  };
  if (!has) {
    for (int i = g.gl; i < nv; i += G) { w.qacc[i] = w.acc_smooth[i]; w.warm[i] = w.acc_smooth[i]; w.frc_con[i] = 0.0; }
    g.sync();
  } else {
    // warm start: keep qacc_warmstart only if it costs less than qacc_smooth
    for (int r = g.gl; r < nefc; r += G) {
      double s1 = 0.0, s2 = 0.0;
      for (int i = 0; i < nv; i++) { s1 += w.J[r * nv + i] * w.warm[i]; s2 += w.J[r * nv + i] * w.acc_smooth[i]; }
      w.jar[r] = s1 - w.efc_aref[r];
      w.jv[r] = s2 - w.efc_aref[r];
    }
    mul_M(m, w, w.Ma, w.warm, g);
    g.sync();
    double cw = constraint_cost(w, w.jar, g);
    double gs = 0.0;
    for (int i = g.gl; i < nv; i += G) gs += (w.Ma[i] - w.frc_smooth[i]) * (w.warm[i] - w.acc_smooth[i]);
    cw += 0.5 * g.sum(gs);
    double cs = constraint_cost(w, w.jv, g);
    bool use_smooth = cw > cs;
    g.sync();
    for (int i = g.gl; i < nv; i += G) w.qacc[i] = use_smooth ? w.acc_smooth[i] : w.warm[i];
    if (use_smooth) {
      for (int r = g.gl; r < nefc; r += G) w.jar[r] = w.jv[r];
      g.sync();
      mul_M(m, w, w.Ma, w.qacc, g);
    }
    g.sync();
  }
  TK_CLK(w, g, 6);
  double scale = 1.0 / (m.meaninertia * (double)max(1, nv));
  // iter == -1 is the initial update(); sharing the call site keeps one copy of the Hessian/Cholesky code and lets
  // all running worlds of the warp execute it together
  int iter = -1;
  bool run = has;
  double alpha = 1.0, oldcost = 0.0;
  while (g.any_block(run)) {
    if (run) {
      if (iter >= 0 && alpha == 0.0) {
        run = false;
      } else {
        if (iter >= 0) {
          for (int i = g.gl; i < nv; i += G) { w.qacc[i] += alpha * w.search[i]; w.Ma[i] += alpha * w.Mv[i]; }
          for (int r = g.gl; r < nefc; r += G) w.jar[r] += alpha * w.jv[r];
          g.sync();
        }
        oldcost = cost;
        TK_CLK(w, g, 12);
        update();
        if (iter >= 0) {
          double gn = 0.0;
          for (int i = g.gl; i < nv; i += G) gn += w.grad[i] * w.grad[i];
          double improvement = scale * (oldcost - cost);
          double gradient = scale * sqrt(g.sum(gn));
          if (improvement < tolerance || gradient < tolerance) run = false;
        }
        iter++;
        if (iter >= iterations) run = false;
      }
    }
    TK_CLK(w, g, 12);
      alpha = primal_search(m, w, g, gauss, tolerance, run);
    TK_CLK(w, g, 11);
  }
  if (has) {
    for (int i = g.gl; i < nv; i += G) w.warm[i] = w.qacc[i];
    g.sync();
  }
}

// mj_EulerSkip + mj_advance, in two parts so that a world which shares its position stage with another one can take
// the factor of M + h D from it (mj_EulerSkip's skipfactor): euler_factor leaves the factor in w.LD / w.Hinv
template <int G, int NV>
TK_DEV void euler_factor(const DM& m, WS& w, const Grp<G, NV>& g) {
  const int nv = NV ? NV : m.nv;
  double h = m.timestep;
  if (m.any_damping) {
    for (int e = g.gl; e < nv * nv; e += G) {
      int i = e / nv;
      w.H[e] = w.M[e] + ((e - i * nv) == i ? h * m.dof_damping[i] : 0.0);
    }
    g.sync();
    factor_ldl(m, w.H, w.LD, w.Hinv, g);
  }
}

template <int G, int NV>
TK_DEV void euler_advance(const DM& m, WS& w, const Grp<G, NV>& g, double& time) {
  const int nv = NV ? NV : m.nv;
  double h = m.timestep;
  if (m.any_damping) {
    for (int i = g.gl; i < nv; i += G) w.tmpv[i] = w.frc_smooth[i] + w.frc_con[i];
    g.sync();
    solve_ldl(m, w.LD, w.Hinv, w.tmpv, g);
  } else {
    for (int i = g.gl; i < nv; i += G) w.tmpv[i] = w.qacc[i];
    g.sync();
  }
  for (int i = g.gl; i < nv; i += G) w.qvel[i] += w.tmpv[i] * h;
  g.sync();
  for (int j = g.gl; j < m.njnt; j += G) {
    int pa = m.jnt_qposadr[j], va = m.jnt_dofadr[j];
    if (m.jnt_type[j] == JNT_FREE) {
      for (int k = 0; k < 3; k++) w.qpos[pa + k] += h * w.qvel[va + k];
      quat_integrate(&w.qpos[pa + 3], &w.qvel[va + 3], h);
    } else {
      w.qpos[pa] += h * w.qvel[va];
    }
  }
  time += h;
  g.sync();
}

template <int G, int NV>
TK_DEV void euler(const DM& m, WS& w, const Grp<G, NV>& g, double& time) {
  euler_factor(m, w, g);
  euler_advance(m, w, g, time);
}

// one full step (mj_step) in two halves, so that a kernel can put a block barrier between them (all warps of an SM
// then run the same region of this large code at the same time and share the instruction cache):
//   step_smooth      position + velocity stages, actuation, smooth acceleration
//   step_constrained constraint solver, acceleration check, Euler
// position stage (mj_fwdPosition): everything that depends on qpos only.  xpos / xquat stay valid until step_velocity.
template <int G, int NV>
TK_DEV void step_position(const DM& m, WS& w, const Grp<G, NV>& g, int& status) {
  for (int i = g.gl; i < m.nq; i += G) if (bad_value(w.qpos[i])) status |= ST_BAD_QPOS;
  for (int i = g.gl; i < m.nv; i += G) if (bad_value(w.qvel[i])) status |= ST_BAD_QVEL;
  kinematics(m, w, g);
  com_pos(m, w, g);
  g.phase();
  TK_CLK(w, g, 0);
  crb_factor(m, w, g);
  TK_CLK(w, g, 15);
  factor_ldl(m, w.M, w.LD, w.LDinv, g);
  g.phase();
  TK_CLK(w, g, 1);
  collision(m, w, g, status);
  g.phase();
  TK_CLK(w, g, 2);
  make_constraint(m, w, g, status);
  g.phase();
  TK_CLK(w, g, 3);
}

// velocity stage + actuation + smooth acceleration (mj_fwdVelocity, mj_fwdActuation, mj_fwdAcceleration)
template <int G, int NV>
TK_DEV void step_velocity(const DM& m, WS& w, const Grp<G, NV>& g) {
  com_vel(m, w, g);
  passive_and_bias(m, w, g);
  g.phase();
  TK_CLK(w, g, 4);
  reference_constraint(m, w, g);
  actuation_acceleration(m, w, g);
  TK_CLK(w, g, 5);
}

// the part of a step after the smooth acceleration, with the Euler factor either computed here (share = nullptr) or
// handed in by the caller through `load_factor` (mj_EulerSkip(skipfactor = 1))
template <int G, int NV>
TK_DEV void check_qacc(const DM& m, const WS& w, const Grp<G, NV>& g, int& status) {
  for (int i = g.gl; i < m.nv; i += G) if (bad_value(w.qacc[i])) status |= ST_BAD_QACC;
}

template <int G, int NV>
TK_DEV void step_smooth(const DM& m, WS& w, const Grp<G, NV>& g, int& status) {
  step_position(m, w, g, status);
  step_velocity(m, w, g);
}

template <int G, int NV>
TK_DEV void step_constrained(const DM& m, WS& w, const Grp<G, NV>& g, int iterations, double tolerance, double& time,
                             int& status) {
  fwd_constraint(m, w, g, iterations, tolerance);
  g.phase();
  for (int i = g.gl; i < m.nv; i += G) if (bad_value(w.qacc[i])) status |= ST_BAD_QACC;
  TK_CLK(w, g, 12);
  euler(m, w, g, time);
  TK_CLK(w, g, 13);
}

// one full step (mj_step).  `iterations`/`tolerance`: solver options of this launch.
template <int G, int NV>
TK_DEV void step(const DM& m, WS& w, const Grp<G, NV>& g, int iterations, double tolerance, double& time, int& status) {
  step_smooth(m, w, g, status);
  g.phase();
  step_constrained(m, w, g, iterations, tolerance, time, status);
}

}  // namespace tk
