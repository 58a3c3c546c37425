// iLQR device kernels: residual FD + Gauss-Newton cost derivatives, Riccati backward pass, acceptance.
// Replaces Differentiator::ResidualDerivatives (src/Differentiator/Differentiator.cpp:464-663),
// Optimiser::ComputeCostDerivatives (src/Optimiser/Optimiser.cpp:197-236), MT::CostDerivativesFromResiduals
// (src/ModelTranslator/ModelTranslator.cpp:552-583), iLQR::BackwardsPassQuuRegularisation
// (src/Optimiser/iLQR.cpp:535-634), Optimiser::SaveBestRollout (Optimiser.cpp:440-468) and
// iLQR::UpdateNominal (iLQR.cpp:936-948).
#include "task_dev.cuh"
#include "traj.h"

namespace tokp {

// one thread per (b, t), t = 0..T: literal residual finite differences
__global__ void k_resid_derivs(DM dm, TaskDev tk_, const double* __restrict__ snap, const double* __restrict__ target,
                               double* __restrict__ rx, double* __restrict__ ru, int B, int T, int SD, int central, double eps,
                               double* __restrict__ scratch, int first, int count) {
  int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= count) return;
  gid += first;  // (b, t) pair first .. first+count-1
  int b = gid / (T + 1);
  int nq = dm.nq, nv = dm.nv, nu = dm.nu, n = tk_.n, m = tk_.m, nr = tk_.nr, nx = 2 * n;
  const double* s = snap + (size_t)gid * SD;
  // private copy of qpos | qvel | ctrl
  double* q = scratch + (size_t)gid * (nq + nv + nu);
  double* v = q + nq;
  double* u = v + nv;
  for (int i = 0; i < nq; i++) q[i] = s[1 + i];
  for (int i = 0; i < nv; i++) v[i] = s[1 + nq + i];
  for (int i = 0; i < nu; i++) u[i] = s[1 + nq + 3 * nv + i];
  const double* tgt = target + (size_t)b * tk_.ntg;
  double r0[kMaxResid], ri[kMaxResid], rd[kMaxResid];
  task_residuals(dm, tk_, q, v, u, tgt, r0);
  double* rxo = rx + (size_t)gid * nr * nx;
  double* ruo = ru + (size_t)gid * nr * m;
  for (int i = 0; i < m; i++) {
    int ua = tk_.ctrl_adr[i];
    double u0 = u[ua];
    bool fwd = !(u0 + eps > tk_.ctrl_limits[2 * i + 1]);
    bool back = (central || !fwd) && !(u0 - eps < tk_.ctrl_limits[2 * i]);
    if (fwd) { u[ua] = u0 + eps; task_residuals(dm, tk_, q, v, u, tgt, ri); }
    if (back) { u[ua] = u0 - eps; task_residuals(dm, tk_, q, v, u, tgt, rd); }
    u[ua] = u0;
    for (int j = 0; j < nr; j++) {
      if (fwd && back) ruo[i * nr + j] = (ri[j] - rd[j]) / (2 * eps);
      else if (fwd) ruo[i * nr + j] = (ri[j] - r0[j]) / eps;
      else if (back) ruo[i * nr + j] = (r0[j] - rd[j]) / eps;
    }
  }
  for (int i = 0; i < n; i++) {
    int va = tk_.dof_adr[i];
    double v0 = v[va];
    v[va] = v0 + eps;
    task_residuals(dm, tk_, q, v, u, tgt, ri);
    if (central) { v[va] = v0 - eps; task_residuals(dm, tk_, q, v, u, tgt, rd); }
    v[va] = v0;
    for (int j = 0; j < nr; j++) rxo[(i + n) * nr + j] = central ? (ri[j] - rd[j]) / (2 * eps) : (ri[j] - r0[j]) / eps;
  }
  for (int i = 0; i < n; i++) {
    // mj_integratePos(qpos, e_dof, +-eps): scalar joints only here (free-joint states: see DESIGN.md)
    int qa = tk_.qpos_adr[i];
    double q0 = q[qa];
    q[qa] = q0 + eps * 1.0;
    task_residuals(dm, tk_, q, v, u, tgt, ri);
    if (central) { q[qa] = q0 + (-eps) * 1.0; task_residuals(dm, tk_, q, v, u, tgt, rd); }
    q[qa] = q0;
    for (int j = 0; j < nr; j++) rxo[i * nr + j] = central ? (ri[j] - rd[j]) / (2 * eps) : (ri[j] - r0[j]) / eps;
  }
}

// one thread per (b, t < T, entry): l_x, l_xx, l_u, l_uu from residuals and their derivatives
__global__ void k_cost_derivs(TaskDev tk_, const double* __restrict__ resid, const double* __restrict__ rx,
                              const double* __restrict__ ru, double* __restrict__ lx, double* __restrict__ lxx,
                              double* __restrict__ lu, double* __restrict__ luu, int B, int T) {
  int n = tk_.n, m = tk_.m, nr = tk_.nr, nx = 2 * n;
  int per = nx + nx * nx + m + m * m;
  long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (long long)B * T * per) return;
  int e = (int)(gid % per);
  long long bt = gid / per;
  int t = (int)(bt % T), b = (int)(bt / T);
  size_t row = (size_t)b * (T + 1) + t;
  const double* w = (t == T - 1) ? tk_.weight_terminal : tk_.weight;  // terminal overwrite at T-1 (Optimiser.cpp:208-211)
  const double* r = resid + row * nr;
  const double* rxo = rx + row * nr * nx;
  const double* ruo = ru + row * nr * m;
  double acc = 0.0;
  if (e < nx) {
    for (int i = 0; i < nr; i++) acc += w[i] * 2 * r[i] * rxo[e * nr + i];
    lx[row * nx + e] = acc;
  } else if (e < nx + nx * nx) {
    int q = e - nx, c = q / nx, a = q - c * nx;
    for (int i = 0; i < nr; i++) acc += w[i] * 2 * rxo[a * nr + i] * rxo[c * nr + i];
    lxx[row * nx * nx + q] = acc;
  } else if (e < nx + nx * nx + m) {
    int a = e - nx - nx * nx;
    for (int i = 0; i < nr; i++) acc += w[i] * 2 * r[i] * ruo[a * nr + i];
    lu[row * m + a] = acc;
  } else {
    int q = e - nx - nx * nx - m, c = q / m, a = q - c * m;
    for (int i = 0; i < nr; i++) acc += 2 * w[i] * ruo[a * nr + i] * ruo[c * nr + i];
    luu[row * m * m + q] = acc;
  }
}

// ---- Riccati backward pass: one CTA per trajectory, serial in t, entries spread over the threads -----------
constexpr int kMaxCtrl = 8;

// Eigen LLT success test (iLQR::CheckMatrixPD)
__device__ bool matrix_pd(const double* A, int n) {
  double L[kMaxCtrl * kMaxCtrl];
  for (int i = 0; i < n * n; i++) L[i] = A[i];
  for (int k = 0; k < n; k++) {
    double x = L[k * n + k];
    for (int j = 0; j < k; j++) x -= L[j * n + k] * L[j * n + k];
    if (x <= 0.0) return false;
    x = sqrt(x);
    L[k * n + k] = x;
    for (int i = k + 1; i < n; i++) {
      double s = L[k * n + i];
      for (int j = 0; j < k; j++) s -= L[j * n + i] * L[j * n + k];
      L[k * n + i] = s / x;
    }
  }
  return true;
}

// Eigen's pivoted LDLT followed by solve(I) (iLQR.cpp:597-600); column-major n x n, n <= kMaxCtrl
// split in two so that the n columns of solve(I) run on n threads: ldlt_factor (one thread, in place in shared
// memory) then ldlt_solve_col (thread c solves column c); per-column arithmetic identical to the serial form
__device__ bool ldlt_factor(double* mat, int n, int* tr) {
#define MM(r, c) mat[(c) * n + (r)]
  bool zero_matrix = false;
  for (int k = 0; k < n; k++) {
    int big = k;
    double bigv = fabs(MM(k, k));
    for (int i = k + 1; i < n; i++)
      if (fabs(MM(i, i)) > bigv) { bigv = fabs(MM(i, i)); big = i; }
    tr[k] = big;
    if (k != big) {
      int s = n - big - 1;
      for (int c = 0; c < k; c++) { double t = MM(k, c); MM(k, c) = MM(big, c); MM(big, c) = t; }
      for (int r = 0; r < s; r++) { double t = MM(big + 1 + r, k); MM(big + 1 + r, k) = MM(big + 1 + r, big); MM(big + 1 + r, big) = t; }
      { double t = MM(k, k); MM(k, k) = MM(big, big); MM(big, big) = t; }
      for (int i = k + 1; i < big; i++) { double t = MM(i, k); MM(i, k) = MM(big, i); MM(big, i) = t; }
    }
    int rs = n - k - 1;
    if (k > 0) {
      double temp[kMaxCtrl];
      for (int j = 0; j < k; j++) temp[j] = MM(j, j) * MM(k, j);
      double s = 0;
      for (int j = 0; j < k; j++) s += MM(k, j) * temp[j];
      MM(k, k) -= s;
      for (int i = 0; i < rs; i++) {
        double t = 0;
        for (int j = 0; j < k; j++) t += MM(k + 1 + i, j) * temp[j];
        MM(k + 1 + i, k) -= t;
      }
    }
    double akk = MM(k, k);
    bool valid = fabs(akk) > 0.0;
    if (k == 0 && !valid) { for (int j = 0; j < n; j++) tr[j] = j; zero_matrix = true; break; }
    if (rs > 0 && valid)
      for (int i = 0; i < rs; i++) MM(k + 1 + i, k) /= akk;
  }
  return zero_matrix;
#undef MM
}

// ldlt_factor with the lanes of one warp on the independent elements of every elimination step (row/column swaps,
// the rows of the trailing column, their division by the pivot); the sums run in the same order per element, so the
// factor is bit-identical to the one-thread form.  Called by all 32 lanes; mat / tr in shared memory.
__device__ bool ldlt_factor_warp(double* mat, int n, int* tr, int lane) {
#define MM(r, c) mat[(c) * n + (r)]
  bool zero_matrix = false;
  for (int k = 0; k < n; k++) {
    int big = k;
    double bigv = fabs(MM(k, k));
    for (int i = k + 1; i < n; i++)
      if (fabs(MM(i, i)) > bigv) { bigv = fabs(MM(i, i)); big = i; }
    __syncwarp();
    if (lane == 0) tr[k] = big;
    if (k != big) {
      // four disjoint element sets (columns < k; columns k / big below row big; the two diagonal entries; the stretch
      // between k and big): no ordering between them
      int s = n - big - 1;
      if (lane < k) { double t = MM(k, lane); MM(k, lane) = MM(big, lane); MM(big, lane) = t; }
      if (lane < s) { double t = MM(big + 1 + lane, k); MM(big + 1 + lane, k) = MM(big + 1 + lane, big); MM(big + 1 + lane, big) = t; }
      if (lane == 31) { double t = MM(k, k); MM(k, k) = MM(big, big); MM(big, big) = t; }
      if (lane > k && lane < big) { double t = MM(lane, k); MM(lane, k) = MM(big, lane); MM(big, lane) = t; }
      __syncwarp();
    }
    int rs = n - k - 1;
    if (k > 0) {
      double temp[kMaxCtrl];
#pragma unroll
      for (int j = 0; j < kMaxCtrl; j++) temp[j] = j < k ? MM(j, j) * MM(k, j) : 0.0;
      if (lane == rs) {  // the diagonal entry, on the first lane that has no row
        double sum = 0;
#pragma unroll
        for (int j = 0; j < kMaxCtrl; j++) if (j < k) sum += MM(k, j) * temp[j];
        MM(k, k) -= sum;
      }
      if (lane < rs) {
        double t = 0;
#pragma unroll
        for (int j = 0; j < kMaxCtrl; j++) if (j < k) t += MM(k + 1 + lane, j) * temp[j];
        MM(k + 1 + lane, k) -= t;
      }
      __syncwarp();
    }
    double akk = MM(k, k);
    bool valid = fabs(akk) > 0.0;
    if (k == 0 && !valid) {
      if (lane < n) tr[lane] = lane;
      zero_matrix = true;
      break;
    }
    if (valid && lane < rs) MM(k + 1 + lane, k) /= akk;
    __syncwarp();
  }
  __syncwarp();
  return zero_matrix;
#undef MM
}

__device__ void ldlt_solve_col(const double* mat, int n, const int* tr, bool zero_matrix, int c, double* Ainv) {
#define MM(r, c) mat[(c) * n + (r)]
  const double tol = 2.2250738585072014e-308;
  {
    double x[kMaxCtrl];
    for (int i = 0; i < n; i++) x[i] = 0.0;
    if (!zero_matrix) {
      x[c] = 1.0;
      for (int k = 0; k < n; k++) { double t = x[k]; x[k] = x[tr[k]]; x[tr[k]] = t; }
      for (int i = 0; i < n; i++)
        for (int j = 0; j < i; j++) x[i] -= MM(i, j) * x[j];
      for (int i = 0; i < n; i++) x[i] = fabs(MM(i, i)) > tol ? x[i] / MM(i, i) : 0.0;
      for (int i = n - 1; i >= 0; i--)
        for (int j = i + 1; j < n; j++) x[i] -= MM(j, i) * x[j];
      for (int k = n - 1; k >= 0; k--) { double t = x[k]; x[k] = x[tr[k]]; x[tr[k]] = t; }
    }
    for (int r = 0; r < n; r++) Ainv[c * n + r] = x[r];
  }
#undef MM
}

__global__ void k_riccati(const double* __restrict__ A, const double* __restrict__ Bm, const double* __restrict__ lx,
                          const double* __restrict__ lxx, const double* __restrict__ lu, const double* __restrict__ luu,
                          double* __restrict__ Kout, double* __restrict__ kout, const double* __restrict__ lambda,
                          int* __restrict__ ok, double* __restrict__ deltaJ, int T, int n, int m) {
  extern __shared__ double sm[];
  int nx = 2 * n, b = blockIdx.x, tid = threadIdx.x, nt = blockDim.x;
  double* Vx = sm;
  double* Vxx = Vx + nx;
  double* Qx = Vxx + nx * nx;
  double* Qu = Qx + nx;
  double* Qxx = Qu + m;
  double* Quu = Qxx + nx * nx;
  double* Qux = Quu + m * m;
  double* VA = Qux + m * nx;
  double* VB = VA + nx * nx;
  double* Qinv = VB + nx * m;
  double* kt = Qinv + m * m;
  double* Kt = kt + m;
  double* tm = Kt + m * nx;    // Quu * k
  double* tmx = tm + m;        // Quu * K
  double* sA = tmx + m * nx;   // A_t, B_t staged per step (coalesced) instead of strided global reads in the products
  double* sB = sA + nx * nx;
  __shared__ int s_fail;
  __shared__ double s_dJ;
  __shared__ double s_mat[kMaxCtrl * kMaxCtrl];
  __shared__ int s_tr[kMaxCtrl];
  const double* Ab = A + (size_t)b * T * nx * nx;
  const double* Bb = Bm + (size_t)b * T * nx * m;
  const double* lxb = lx + (size_t)b * (T + 1) * nx;
  const double* lxxb = lxx + (size_t)b * (T + 1) * nx * nx;
  const double* lub = lu + (size_t)b * (T + 1) * m;
  const double* luub = luu + (size_t)b * (T + 1) * m * m;
  double lam = lambda[b];
  if (lam < 0.0) return;  // trajectory not part of this pass
  for (int i = tid; i < nx; i += nt) Vx[i] = lxb[(size_t)(T - 1) * nx + i];
  for (int i = tid; i < nx * nx; i += nt) Vxx[i] = lxxb[(size_t)(T - 1) * nx * nx + i];
  if (tid == 0) { s_fail = 0; s_dJ = 0.0; }
  __syncthreads();
  int pd_counter = 0;
  for (int t = T - 1; t >= 0; t--) {
    pd_counter++;
    for (int i = tid; i < nx * nx; i += nt) sA[i] = Ab[(size_t)t * nx * nx + i];
    for (int i = tid; i < nx * m; i += nt) sB[i] = Bb[(size_t)t * nx * m + i];
    __syncthreads();
    const double* At = sA;
    const double* Bt = sB;
    for (int i = tid; i < nx + m; i += nt) {
      double s = 0;
      if (i < nx) { for (int r = 0; r < nx; r++) s += At[i * nx + r] * Vx[r]; Qx[i] = lxb[(size_t)t * nx + i] + s; }
      else { int c = i - nx; for (int r = 0; r < nx; r++) s += Bt[c * nx + r] * Vx[r]; Qu[c] = lub[(size_t)t * m + c] + s; }
    }
    for (int e = tid; e < nx * nx + nx * m; e += nt) {
      double s = 0;
      if (e < nx * nx) { int c = e / nx, r = e - c * nx; for (int q = 0; q < nx; q++) s += Vxx[q * nx + r] * At[c * nx + q]; VA[e] = s; }
      else { int e2 = e - nx * nx, c = e2 / nx, r = e2 - c * nx; for (int q = 0; q < nx; q++) s += Vxx[q * nx + r] * Bt[c * nx + q]; VB[e2] = s; }
    }
    __syncthreads();
    // Warp 0 takes the small-matrix chain (Q_uu, its regularised pivoted LDLT, inv(Q_uu) by m column solves, k, Q_uu k,
    // delta_J - mostly single-lane work) while the other warps compute the two large products Q_xx = l_xx + A' (V_xx A)
    // and Q_ux = B' (V_xx A): the serial chain hides behind the 18 x 18 x 18 product instead of standing between
    // block barriers.  Same arithmetic per entry as the straight sequence.
    if (tid < 32) {
      const int lane = tid;
      for (int e2 = lane; e2 < m * m; e2 += 32) {
        int c = e2 / m, r = e2 - c * m;
        double s = 0;
        for (int q = 0; q < nx; q++) s += Bt[r * nx + q] * VB[c * nx + q];
        Quu[e2] = luub[(size_t)t * m * m + e2] + s;
      }
      __syncwarp();
      for (int i = lane; i < m * m; i += 32) {
        double v = Quu[i];
        if ((i / m) == (i - (i / m) * m)) v += lam;
        s_mat[i] = v;
      }
      __syncwarp();
      if (pd_counter >= 100) {
        if (lane == 0 && !matrix_pd(s_mat, m)) s_fail = 1;
        __syncwarp();
      }
      if (!s_fail) {
        bool zero = ldlt_factor_warp(s_mat, m, s_tr, lane);
        if (lane < m) ldlt_solve_col(s_mat, m, s_tr, zero, lane, Qinv);
        __syncwarp();
        if (lane < m) {
          double sum = 0;
          for (int q = 0; q < m; q++) sum += Qinv[q * m + lane] * Qu[q];
          kt[lane] = -sum;
        }
        __syncwarp();
        if (lane < m) {
          double sum = 0;
          for (int q = 0; q < m; q++) sum += Quu[q * m + lane] * kt[q];
          tm[lane] = sum;
        }
        __syncwarp();
        if (lane == 0) {
          double d1 = 0, d2 = 0;
          for (int r = 0; r < m; r++) { d1 += kt[r] * Qu[r]; d2 += kt[r] * tm[r]; }
          s_dJ += d1;
          s_dJ += d2;
        }
      }
    } else {
      for (int e = tid - 32; e < nx * nx + m * nx; e += nt - 32) {
        double s = 0;
        if (e < nx * nx) {
          int c = e / nx, r = e - c * nx;
          for (int q = 0; q < nx; q++) s += At[r * nx + q] * VA[c * nx + q];
          Qxx[e] = lxxb[(size_t)t * nx * nx + e] + s;
        } else {
          int e2 = e - nx * nx, c = e2 / m, r = e2 - c * m;
          for (int q = 0; q < nx; q++) s += Bt[r * nx + q] * VA[c * nx + q];
          Qux[e2] = s;
        }
      }
    }
    if (pd_counter >= 100) pd_counter = 0;
    __syncthreads();
    if (s_fail) break;
    for (int e = tid; e < m * nx; e += nt) {
      int c = e / m, r = e - c * m;
      double s = 0;
      for (int q = 0; q < m; q++) s += Qinv[q * m + r] * Qux[c * m + q];
      Kt[e] = -s;
    }
    __syncthreads();
    for (int e = tid; e < m * nx; e += nt) {
      int c = e / m, r = e - c * m;
      double s = 0;
      for (int q = 0; q < m; q++) s += Quu[q * m + r] * Kt[c * m + q];
      tmx[e] = s;
    }
    __syncthreads();
    for (int i = tid; i < nx; i += nt) {
      double a = 0, bb = 0, c = 0;
      for (int r = 0; r < m; r++) { a += Kt[i * m + r] * tm[r]; bb += Kt[i * m + r] * Qu[r]; c += Qux[i * m + r] * kt[r]; }
      Vx[i] = Qx[i] + a + bb + c;
    }
    for (int e = tid; e < nx * nx; e += nt) {
      int c = e / nx, r = e - c * nx;
      double a = 0, bb = 0, d = 0;
      for (int q = 0; q < m; q++) { a += Kt[r * m + q] * tmx[c * m + q]; bb += Kt[r * m + q] * Qux[c * m + q]; d += Qux[r * m + q] * Kt[c * m + q]; }
      VA[e] = Qxx[e] + a + bb + d;  // VA reused as the unsymmetrised V_xx
    }
    for (int e = tid; e < m * nx + m; e += nt) {
      if (e < m * nx) Kout[((size_t)b * T + t) * m * nx + e] = Kt[e];
      else kout[((size_t)b * T + t) * m + (e - m * nx)] = kt[e - m * nx];
    }
    __syncthreads();
    for (int e = tid; e < nx * nx; e += nt) {
      int c = e / nx, r = e - c * nx;
      Vxx[e] = (VA[c * nx + r] + VA[r * nx + c]) / 2;
    }
    __syncthreads();
  }
  if (tid == 0) { ok[b] = s_fail ? 0 : 1; deltaJ[b] = s_dJ; }
}

// SaveBestRollout + UpdateNominal for accepted trajectories: one thread per (b, t' = 0..T)
__global__ void k_accept(DM dm, TaskDev tk_, double* __restrict__ snap, const double* __restrict__ cand,
                         double* __restrict__ X, double* __restrict__ U, double* __restrict__ resid,
                         const double* __restrict__ target, const int* __restrict__ accept, int B, int T, int SD, int NA) {
  int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= B * (T + 1)) return;
  int b = gid / (T + 1), t = gid - b * (T + 1);
  int ai = accept[b];
  if (ai < 0) return;
  int nq = dm.nq, nv = dm.nv, nu = dm.nu, n = tk_.n, m = tk_.m, nx = 2 * n, nr = tk_.nr;
  int so = 1 + nq + 3 * nv;  // offset of ctrl in the device snapshot
  double* s = snap + ((size_t)b * (T + 1) + t) * SD;
  const double* c = cand + (((size_t)b * NA + ai) * (T + 1) + t) * SD;
  if (t >= 1) {
    // saved[t] <- rollout_data[t] (time, qpos, qvel, qacc, warmstart); its ctrl is still the old one here
    for (int i = 0; i < so; i++) s[i] = c[i];
    double r[kMaxResid];
    task_residuals(dm, tk_, s + 1, s + 1 + nq, s + so, target + (size_t)b * tk_.ntg, r);
    for (int i = 0; i < nr; i++) resid[((size_t)b * (T + 1) + t) * nr + i] = r[i];
    double* Xo = X + ((size_t)b * (T + 1) + t) * nx;
    for (int i = 0; i < n; i++) { Xo[i] = s[1 + tk_.qpos_adr[i]]; Xo[n + i] = s[1 + nq + tk_.dof_adr[i]]; }
  }
  if (t < T) {
    for (int i = 0; i < nu; i++) s[so + i] = c[so + i];
    for (int i = 0; i < m; i++) U[((size_t)b * T + t) * m + i] = s[so + tk_.ctrl_adr[i]];
  }
}

}  // namespace tokp

// ================================= host side of the C ABI =================================================
namespace tokp {
int launch_rollout(tokp_traj* tr, int mode, const double* d_Uin, const double* d_alphas, int nalpha);
int generate_keypoints(tokp_traj* tr);
int interpolate(tokp_traj* tr);
int launch_fd(tokp_traj* tr, const unsigned long long* mask, const int* rows, int nrows, int central, double eps);
void pack_snapshot(const tokp_traj* tr, const double* s, double* d);
int check_applied_zero(const tokp_traj* tr, const double* s);
int read_keypoints(tokp_traj* tr, int* counts, int* indices, double* pct);
int update_fd_counters(tokp_traj* tr);
int kp_stats_enqueue(tokp_traj* tr);
int kp_stats_finish(tokp_traj* tr, bool set_counters, const char* what);
int check_world_status(tokp_traj* tr, int column, const char* what);

static int cost_derivs(tokp_traj* tr, int central, double eps) {
  int B = tr->B, T = tr->T;
  if (tr->nr > kMaxResid) { set_error("too many residuals"); return TOKP_ERR_UNSUPPORTED; }
  k_resid_derivs<<<(B * (T + 1) + 63) / 64, 64, 0, tr->stream>>>(tr->dm, tr->task, tr->snap, tr->target, tr->rx, tr->ru, B, T,
                                                                 tr->SD, central, eps, tr->fd_out, 0, B * (T + 1));
  int per = tr->nx + tr->nx * tr->nx + tr->m + tr->m * tr->m;
  long long tot = (long long)B * T * per;
  k_cost_derivs<<<(int)((tot + 255) / 256), 256, 0, tr->stream>>>(tr->task, tr->resid, tr->rx, tr->ru, tr->lx, tr->lxx, tr->lu,
                                                                    tr->luu, B, T);
  g_launches += 2;
  TOKP_CUDA(cudaGetLastError());
  return TOKP_OK;
}

// Optimiser::FilterDynamicsMatrices (Optimiser.cpp:340-406): one thread per (b, velocity row, column) walks the entry's
// signal over t = 0..T-1.  low_pass is a first-order recursion (sequential in t); FIR is a short causal convolution,
// computed back to front so that it can run in place.
struct FilterArgs {
  int method, nfir;
  double a;
  double fir[16];
};

__global__ void k_filter(double* __restrict__ A, int B, int T, int n, FilterArgs f) {
  int nx = 2 * n;
  int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= B * n * nx) return;
  int b = gid / (n * nx);
  int e = gid - b * n * nx;
  int col = e / n, row = n + (e - col * n);
  double* x = A + (size_t)b * T * nx * nx + (size_t)col * nx + row;
  const size_t st = (size_t)nx * nx;
  if (f.method == TOKP_FILTER_LOW_PASS) {
    double yprev = x[0], xprev = x[0];
    for (int t = 0; t < T; t++) {
      double xt = x[t * st];
      double y = __dadd_rn(__dmul_rn(1 - f.a, yprev), __dmul_rn(f.a, __dadd_rn(xt, xprev) / 2));
      x[t * st] = y;
      xprev = xt;
      yprev = y;
    }
  } else {
    for (int t = T - 1; t >= 0; t--) {
      double acc = 0;
      for (int j = 0; j < f.nfir && j <= t; j++) acc = __dadd_rn(acc, __dmul_rn(x[(t - j) * st], f.fir[j]));
      x[t * st] = acc;
    }
  }
}

static int filter_dynamics(tokp_traj* tr) {
  if (tr->filter_method == TOKP_FILTER_NONE) return TOKP_OK;
  FilterArgs f;
  f.method = tr->filter_method; f.nfir = tr->nfir; f.a = tr->low_pass_a;
  for (int i = 0; i < 16; i++) f.fir[i] = tr->fir[i];
  int tot = tr->B * tr->n * tr->nx;
  k_filter<<<(tot + 127) / 128, 128, 0, tr->stream>>>(tr->A, tr->B, tr->T, tr->n, f);
  g_launches++;
  TOKP_CUDA(cudaGetLastError());
  return TOKP_OK;
}

static int generate_derivatives(tokp_traj* tr) {
  int rc = generate_keypoints(tr);
  if (rc) return rc;
  if (tr->task.kp_method != TOKP_KP_ITERATIVE_ERROR) {
    rc = launch_fd(tr, tr->kp_mask, nullptr, tr->B * tr->T, 1, 1e-6);
    if (rc) return rc;
  }
  rc = kp_stats_enqueue(tr);  // finished by the caller once it has synchronised (kp_stats_finish)
  if (rc) return rc;
  rc = interpolate(tr);
  if (rc) return rc;
  rc = cost_derivs(tr, 1, 1e-6);
  if (rc) return rc;
  return filter_dynamics(tr);  // Optimiser.cpp:104-107: after the cost derivatives, before the backward pass
}

static int backward_pass(tokp_traj* tr) {
  int nx = tr->nx, m = tr->m;
  if (m > kMaxCtrl || m < 1) { set_error("backward pass supports 1..8 controls"); return TOKP_ERR_UNSUPPORTED; }
  size_t smem = (size_t)(4 * nx * nx + 2 * nx + 5 * m * nx + 2 * m * m + 3 * m) * 8;
  if (smem > 227 * 1024) { set_error("state dimension too large for the backward-pass workspace"); return TOKP_ERR_UNSUPPORTED; }
  TOKP_CUDA(cudaFuncSetAttribute(k_riccati, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int threads = nx * nx >= 1024 ? 256 : 128;
  k_riccati<<<tr->B, threads, smem, tr->stream>>>(tr->A, tr->Bm, tr->lx, tr->lxx, tr->lu, tr->luu, tr->K, tr->k, tr->lambda_d,
                                                  tr->bp_ok, tr->deltaJ, tr->T, tr->n, m);
  g_launches++;
  TOKP_CUDA(cudaGetLastError());
  return TOKP_OK;
}

static int accept(tokp_traj* tr, const std::vector<int>& idx) {
  int B = tr->B, T = tr->T;
  int* d_idx = reinterpret_cast<int*>(tr->status);  // scratch: status is only live during the FD phase
  TOKP_CUDA(cudaMemcpyAsync(d_idx, idx.data(), B * 4, cudaMemcpyHostToDevice, tr->stream));
  k_accept<<<(B * (T + 1) + 127) / 128, 128, 0, tr->stream>>>(tr->dm, tr->task, tr->snap, tr->cand, tr->X, tr->U, tr->resid,
                                                             tr->target, d_idx, B, T, tr->SD, tokp_traj::NA);
  g_launches++;
  TOKP_CUDA(cudaGetLastError());
  return TOKP_OK;
}

struct Timer {
  tokp_traj* tr;
  explicit Timer(tokp_traj* t) : tr(t) { cudaEventRecord(tr->ev0, tr->stream); }
  double stop() {
    cudaEventRecord(tr->ev1, tr->stream);
    cudaEventSynchronize(tr->ev1);
    float ms = 0;
    cudaEventElapsedTime(&ms, tr->ev0, tr->ev1);
    return ms;
  }
};
}  // namespace tokp

using namespace tokp;

// scatter the resident start snapshots into saved[b][0]
__global__ void k_seed_snap0(double* __restrict__ snap, const double* __restrict__ snap0, int B, int T, int SD) {
  int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= B * SD) return;
  int b = gid / SD, i = gid - b * SD;
  snap[(size_t)b * (T + 1) * SD + i] = snap0[gid];
}

extern "C" int tokp_rollout(tokp_traj* tr, const double* snapshot0, const double* U, double* cost) {
  if (!tr || (!snapshot0) != (!U)) { set_error("tokp_rollout: bad argument"); return TOKP_ERR_ARG; }
  int B = tr->B, T = tr->T, m = tr->m;
  if (snapshot0) {
    std::vector<double> d((size_t)B * tr->SD);
    for (int b = 0; b < B; b++) {
      int rc = check_applied_zero(tr, snapshot0 + (size_t)b * tr->S);
      if (rc) return rc;
      pack_snapshot(tr, snapshot0 + (size_t)b * tr->S, d.data() + (size_t)b * tr->SD);
    }
    TOKP_CUDA(cudaMemcpyAsync(tr->snap0_in, d.data(), d.size() * 8, cudaMemcpyHostToDevice, tr->stream));
    TOKP_CUDA(cudaMemcpyAsync(tr->U0_in, U, (size_t)B * T * m * 8, cudaMemcpyHostToDevice, tr->stream));
    TOKP_CUDA(cudaStreamSynchronize(tr->stream));
    tr->have_inputs = true;
  } else if (!tr->have_inputs) {
    set_error("tokp_rollout: no resident inputs (call once with host buffers first)");
    return TOKP_ERR_STATE;
  }
  k_seed_snap0<<<(B * tr->SD + 127) / 128, 128, 0, tr->stream>>>(tr->snap, tr->snap0_in, B, T, tr->SD);
  g_launches++;
  int rc = launch_rollout(tr, 0, tr->U0_in, nullptr, 1);
  if (rc) return rc;
  std::vector<double> hc(B);
  TOKP_CUDA(cudaMemcpyAsync(hc.data(), tr->cost, B * 8, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  if (cost) for (int b = 0; b < B; b++) cost[b] = hc[b];
  tr->have_rollout = true;
  return TOKP_OK;
}

extern "C" int tokp_cost_derivs(tokp_traj* tr, int central_diff, double eps) {
  if (!tr || !(eps > 0)) { set_error("tokp_cost_derivs: bad argument"); return TOKP_ERR_ARG; }
  if (!tr->have_rollout) { set_error("tokp_cost_derivs: no nominal trajectory yet"); return TOKP_ERR_STATE; }
  int rc = cost_derivs(tr, central_diff, eps);
  if (rc) return rc;
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  return TOKP_OK;
}

extern "C" int tokp_traj_set_filter(tokp_traj* tr, int method, double low_pass_a, const double* fir, int nfir) {
  if (!tr || method < TOKP_FILTER_NONE || method > TOKP_FILTER_FIR || nfir < 0 || nfir > 16) {
    set_error("tokp_traj_set_filter: bad argument (method 0..2, at most 16 FIR taps)");
    return TOKP_ERR_ARG;
  }
  tr->filter_method = method;
  tr->low_pass_a = low_pass_a;
  if (fir && nfir > 0) { tr->nfir = nfir; for (int i = 0; i < nfir; i++) tr->fir[i] = fir[i]; }
  return TOKP_OK;
}

extern "C" int tokp_filter_dynamics(tokp_traj* tr) {
  if (!tr) { set_error("tokp_filter_dynamics: null trajectory"); return TOKP_ERR_ARG; }
  if (tr->filter_method == TOKP_FILTER_NONE) { set_error("tokp_filter_dynamics: filtering method not recognised (none set)"); return TOKP_ERR_STATE; }
  int rc = filter_dynamics(tr);
  if (rc) return rc;
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  return TOKP_OK;
}

extern "C" int tokp_residual_derivs(tokp_traj* tr, int b, int t, int central_diff, double eps, double* r_x, double* r_u) {
  if (!tr || b < 0 || b >= tr->B || t < 0 || t > tr->T || !(eps > 0) || !r_x) { set_error("tokp_residual_derivs: bad argument"); return TOKP_ERR_ARG; }
  if (tr->nr > kMaxResid) { set_error("too many residuals"); return TOKP_ERR_UNSUPPORTED; }
  int first = b * (tr->T + 1) + t;
  k_resid_derivs<<<1, 64, 0, tr->stream>>>(tr->dm, tr->task, tr->snap, tr->target, tr->rx, tr->ru, tr->B, tr->T, tr->SD,
                                           central_diff, eps, tr->fd_out, first, 1);
  g_launches++;
  TOKP_CUDA(cudaGetLastError());
  TOKP_CUDA(cudaMemcpyAsync(r_x, tr->rx + (size_t)first * tr->nr * tr->nx, (size_t)tr->nr * tr->nx * 8, cudaMemcpyDeviceToHost, tr->stream));
  if (r_u && tr->m)
    TOKP_CUDA(cudaMemcpyAsync(r_u, tr->ru + (size_t)first * tr->nr * tr->m, (size_t)tr->nr * tr->m * 8, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  return TOKP_OK;
}

extern "C" int tokp_generate_derivatives(tokp_traj* tr) {
  if (!tr) { set_error("tokp_generate_derivatives: null trajectory"); return TOKP_ERR_ARG; }
  if (!tr->have_rollout) { set_error("tokp_generate_derivatives: no nominal trajectory yet"); return TOKP_ERR_STATE; }
  Timer tm(tr);
  int rc = generate_derivatives(tr);
  if (rc) return rc;
  tr->timers[0] = tm.stop();
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  // the iterative_error level loop keeps its own world / column counters
  return kp_stats_finish(tr, tr->task.kp_method != TOKP_KP_ITERATIVE_ERROR, "tokp_generate_derivatives");
}

extern "C" int tokp_backward_pass(tokp_traj* tr, const double* lambda, int* ok, double* delta_J) {
  if (!tr || !lambda) { set_error("tokp_backward_pass: bad argument"); return TOKP_ERR_ARG; }
  int B = tr->B;
  TOKP_CUDA(cudaMemcpyAsync(tr->lambda_d, lambda, B * 8, cudaMemcpyHostToDevice, tr->stream));
  Timer tm(tr);
  int rc = backward_pass(tr);
  if (rc) return rc;
  tr->timers[1] = tm.stop();
  std::vector<int> hok(B);
  std::vector<double> hdj(B);
  TOKP_CUDA(cudaMemcpyAsync(hok.data(), tr->bp_ok, B * 4, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaMemcpyAsync(hdj.data(), tr->deltaJ, B * 8, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  for (int b = 0; b < B; b++) { if (ok) ok[b] = hok[b]; if (delta_J) delta_J[b] = hdj[b]; }
  return TOKP_OK;
}

extern "C" int tokp_forward_pass(tokp_traj* tr, const double* alphas, int nalpha, double* costs) {
  if (!tr || !alphas || nalpha < 1 || nalpha > tokp_traj::NA) { set_error("tokp_forward_pass: bad argument"); return TOKP_ERR_ARG; }
  int B = tr->B, NA = tokp_traj::NA;
  double ha[tokp_traj::NA];
  for (int i = 0; i < NA; i++) ha[i] = alphas[i < nalpha ? i : nalpha - 1];
  double* d_al = tr->fd_out;  // staging (free outside the FD phase)
  TOKP_CUDA(cudaMemcpyAsync(d_al, ha, NA * 8, cudaMemcpyHostToDevice, tr->stream));
  Timer tm(tr);
  int rc = launch_rollout(tr, 1, nullptr, d_al, nalpha);
  if (rc) return rc;
  tr->timers[2] = tm.stop();
  std::vector<double> hc((size_t)B * NA);
  TOKP_CUDA(cudaMemcpyAsync(hc.data(), tr->cand_cost, hc.size() * 8, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  if (costs)
    for (int b = 0; b < B; b++)
      for (int i = 0; i < nalpha; i++) costs[(size_t)b * nalpha + i] = hc[(size_t)b * NA + i];
  return TOKP_OK;
}

extern "C" int tokp_accept(tokp_traj* tr, int b, int alpha_idx) {
  if (!tr || b < 0 || b >= tr->B || alpha_idx < 0 || alpha_idx >= tokp_traj::NA) { set_error("tokp_accept: bad argument"); return TOKP_ERR_ARG; }
  std::vector<int> idx(tr->B, -1);
  idx[b] = alpha_idx;
  int rc = accept(tr, idx);
  if (rc) return rc;
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  return TOKP_OK;
}

// iLQR::Optimise + iLQR::Iteration for B trajectories in lock-step (per-trajectory lambda / convergence state)
extern "C" int tokp_optimise(tokp_traj* tr, const double* snapshot0, const double* U_init, int max_iterations,
                             int min_iterations, const double* lambda_init, double* U_out, double* stats) {
  // snapshot0 == U_init == NULL: re-use the device-resident inputs of the previous call (no bulk H2D)
  if (!tr || (!snapshot0) != (!U_init) || max_iterations < 0) { set_error("tokp_optimise: bad argument"); return TOKP_ERR_ARG; }
  int B = tr->B, T = tr->T, m = tr->m, NA = tokp_traj::NA;
  const double max_lambda = 10.0, min_lambda = 0.0001, lambda_factor = 10.0, eps_converge = 0.02;
  std::vector<double> old_cost(B), new_cost(B), initial(B), lambda(B, 0.1), t_der(B, 0), t_bp(B, 0), t_fp(B, 0), pct_sum(B, 0);
  std::vector<int> done(B, 0), reduced(B, 1), iters(B, 0), pct_cnt(B, 0);
  if (lambda_init) for (int b = 0; b < B; b++) lambda[b] = lambda_init[b];
  int rc = tokp_rollout(tr, snapshot0, U_init, old_cost.data());
  if (rc) return rc;
  new_cost = old_cost;
  initial = old_cost;
  double alphas[tokp_traj::NA];
  for (int i = 1; i < NA + 1; i++) { double lin = (double)i / NA; alphas[i - 1] = lin * lin; }
  std::vector<double> costs((size_t)B * NA), dJ(B);
  std::vector<int> ok(B);
  for (int it = 0; it < max_iterations; it++) {
    bool any_active = false;
    for (int b = 0; b < B; b++) if (!done[b]) { any_active = true; iters[b]++; }
    if (!any_active) break;
    // STEP 1: derivatives (only trajectories whose cost reduced last iteration need them; recomputing for the
    // others is idempotent because their nominal trajectory did not change)
    bool need = false;
    for (int b = 0; b < B; b++) if (!done[b] && reduced[b]) need = true;
    if (need) {
      rc = tokp_generate_derivatives(tr);  // also checks the status words of the FD worlds
      if (rc) return rc;
      for (int b = 0; b < B; b++)
        if (!done[b] && reduced[b]) {
          const double s = tr->h_kp_stats[3 * b];  // sum over the dofs of the key-point percentages (k_kp_stats)
          pct_sum[b] += s / tr->n; pct_cnt[b]++;
          t_der[b] += tr->timers[0];
        }
    }
    // STEP 2: backward pass with lambda escalation
    std::vector<int> valid(B, 0), lambda_exit(B, 0);
    std::vector<double> lam_pass(B);
    for (;;) {
      bool pending = false;
      for (int b = 0; b < B; b++) {
        bool p = !done[b] && !valid[b] && !lambda_exit[b];
        lam_pass[b] = p ? lambda[b] : -1.0;  // negative lambda: the kernel leaves this trajectory untouched
        pending = pending || p;
      }
      if (!pending) break;
      rc = tokp_backward_pass(tr, lam_pass.data(), ok.data(), dJ.data());
      if (rc) return rc;
      for (int b = 0; b < B; b++) {
        if (lam_pass[b] < 0) continue;
        t_bp[b] += tr->timers[1];
        valid[b] = ok[b];
        if (!ok[b]) lambda[b] *= lambda_factor; else lambda[b] /= lambda_factor;
        if (lambda[b] > max_lambda) { lambda[b] = max_lambda; lambda_exit[b] = 1; }
        if (lambda[b] < min_lambda) lambda[b] = min_lambda;
      }
    }
    for (int b = 0; b < B; b++) if (!done[b] && lambda_exit[b]) done[b] = 1;
    // STEP 3: alpha-batched forward passes
    rc = tokp_forward_pass(tr, alphas, NA, costs.data());
    if (rc) return rc;
    std::vector<int> acc(B, -1);
    for (int b = 0; b < B; b++) {
      if (done[b]) continue;
      t_fp[b] += tr->timers[2];
      int best = 0;
      for (int i = 1; i < NA; i++) if (costs[(size_t)b * NA + i] < costs[(size_t)b * NA + best]) best = i;
      if (costs[(size_t)b * NA + best] < old_cost[b]) { new_cost[b] = costs[(size_t)b * NA + best]; acc[b] = best; }
      else new_cost[b] = old_cost[b];
    }
    rc = accept(tr, acc);
    if (rc) return rc;
    TOKP_CUDA(cudaStreamSynchronize(tr->stream));
    // STEP 4: convergence / lambda bookkeeping
    for (int b = 0; b < B; b++) {
      if (done[b]) continue;
      bool converged = ((old_cost[b] - new_cost[b]) / new_cost[b]) < eps_converge;
      if (new_cost[b] < old_cost[b]) { old_cost[b] = new_cost[b]; reduced[b] = 1; }
      else {
        reduced[b] = 0;
        lambda[b] *= lambda_factor; lambda[b] *= lambda_factor;
        if (lambda[b] > max_lambda) lambda[b] = max_lambda;
      }
      if (converged && it >= min_iterations) done[b] = 1;
    }
  }
  rc = check_world_status(tr, 1, "tokp_optimise");  // unsupported features inside a rollout
  if (rc) return rc;
  if (U_out) TOKP_CUDA(cudaMemcpy(U_out, tr->U, (size_t)B * T * m * 8, cudaMemcpyDeviceToHost));
  if (stats)
    for (int b = 0; b < B; b++) {
      double* s = stats + (size_t)b * 8;
      s[0] = initial[b]; s[1] = new_cost[b]; s[2] = iters[b]; s[3] = lambda[b];
      s[4] = pct_cnt[b] ? pct_sum[b] / pct_cnt[b] : 0.0; s[5] = t_der[b]; s[6] = t_bp[b]; s[7] = t_fp[b];
    }
  return TOKP_OK;
}
