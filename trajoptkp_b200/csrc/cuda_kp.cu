// Key-point selection and derivative interpolation on the device.
// Replaces KeypointGenerator::GenerateKeyPoints (src/KeyPointGenerator/KeyPointGenerator.cpp:76-135):
// set_interval :319-339, adaptive (jerk/accel) :341-382 with profiles :730-795, velocity_change :642-728,
// iterative_error :449-640 (level-synchronous batches), percentages :814-838, and
// InterpolateDerivatives :840-954.  Rows are kept as 64-bit DoF masks; the integer outputs are bit-exact
// because every comparison uses the same IEEE operations in the same order (no contraction possible).
#include <algorithm>

#include "traj.h"

namespace tokp {

int launch_fd(tokp_traj* tr, const unsigned long long* mask, const int* rows, int nrows, int central, double eps);
void adapt_fast_capacity(tokp_traj* tr, long long worlds, long long overflowed);

__global__ void k_kp_set_interval(unsigned long long* mask, int B, int T, int n, int min_N) {
  int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= B * T) return;
  int t = gid % T;
  unsigned long long full = n >= 64 ? ~0ull : ((1ull << n) - 1);
  mask[gid] = (t == T - 1 || t % min_N == 0) ? full : 0ull;
}

// one thread per (b, dof): serial scan over t.  mode 0: jerk profile, 1: acceleration profile, 2: velocity change
__global__ void k_kp_scan(unsigned long long* mask, unsigned long long* last_row, const double* __restrict__ X, int B, int T, int n, int mode, int min_N,
                          int max_N, const double* __restrict__ thr, double dt) {
  int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= B * n) return;
  int b = gid / n, j = gid - b * n;
  int nx = 2 * n;
  const double* Xb = X + (size_t)b * (T + 1) * nx;
  unsigned long long* mb = mask + (size_t)b * T;
  unsigned long long bit = 1ull << j;
  auto vel = [&](int t) { return Xb[(size_t)t * nx + n + j]; };
  atomicOr(&mb[0], bit);
  if (mode < 2) {
    int last = 0;
    double th = thr[j];
    for (int t = 1; t < T - 1; t++) {
      double p;
      if (mode == 0) {
        if (t < T - 2) {
          double a1 = (vel(t + 1) - vel(t)) / dt;
          double a2 = (vel(t + 2) - vel(t + 1)) / dt;
          p = fabs((a2 - a1) / dt);
        } else {
          p = 0.0;
        }
      } else {
        p = vel(t + 1) - vel(t);
      }
      bool hit = false;
      if ((t - last) >= min_N) {
        if (p > th) { hit = true; last = t; }
      }
      if ((t - last) >= max_N) { hit = true; last = t; }
      if (hit) atomicOr(&mb[t], bit);
    }
    atomicOr(&mb[T - 1], bit);
  } else {
    int counter = 0;
    double last_val = 0.0, last_dir = 0.0, th = thr[j];
    for (int t = 1; t < T; t++) {
      counter++;
      double dir = vel(t) - vel(t - 1);
      last_val += fabs(vel(t));
      bool hit = false;
      if (counter >= min_N && fabs(last_val) > th) hit = true;
      if (!hit) {
        if (counter >= min_N) {
          if (dir * last_dir < 0) hit = true;
        } else {
          last_dir = dir;
        }
      }
      if (!hit && counter >= max_N) hit = true;
      if (hit) { atomicOr(&mb[t], bit); last_val = 0.0; counter = 0; if (t == T - 1) atomicOr(&last_row[b], bit); }
    }
    atomicOr(&mb[T - 1], bit);
  }
}

// per (b, dof): previous / next key-point index at every t (next == t when t itself is a key-point)
__global__ void k_kp_links(const unsigned long long* __restrict__ mask, int* __restrict__ prev, int* __restrict__ next, int B,
                           int T, int n) {
  int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= B * n) return;
  int b = gid / n, j = gid - b * n;
  const unsigned long long* mb = mask + (size_t)b * T;
  int* pv = prev + (size_t)b * T * n;
  int* nx_ = next + (size_t)b * T * n;
  // the reference preloads startIndices = 0 for every dof (KeyPointGenerator.cpp:871-875)
  int last = 0;
  for (int t = 0; t < T; t++) {
    pv[(size_t)t * n + j] = last;
    if (t > 0 && ((mb[t] >> j) & 1ull)) last = t;
  }
  int nxt = -1;
  for (int t = T - 1; t >= 0; t--) {
    if (t > 0 && ((mb[t] >> j) & 1ull)) nxt = t;
    nx_[(size_t)t * n + j] = nxt;
  }
}

// col[k] = col[s] + (k - s) * ((col[e] - col[s]) / (e - s)), one thread per (b, t, dof, row)
__global__ void k_interp(double* __restrict__ A, double* __restrict__ Bm, const unsigned long long* __restrict__ mask,
                         const int* __restrict__ prev, const int* __restrict__ next, int B, int T, int n, int m) {
  int nx = 2 * n;
  long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long total = (long long)B * T * n * nx;
  if (gid >= total) return;
  int r = (int)(gid % nx);
  long long q = gid / nx;
  int i = (int)(q % n);
  q /= n;
  int t = (int)(q % T);
  int b = (int)(q / T);
  if (t == 0) return;
  size_t row = (size_t)b * T + t;
  if ((mask[row] >> i) & 1ull) return;  // key-point rows keep their computed columns
  int s = prev[row * n + i], e = next[row * n + i];
  if (e < 0) return;  // no later key-point of this dof: untouched, like the reference
  double denom = (double)(e - s), f = (double)(t - s);
  double* Ab = A + (size_t)b * T * nx * nx;
  auto Ael = [&](int tt, int c) -> double& { return Ab[(size_t)tt * nx * nx + (size_t)c * nx + r]; };
  double a0 = Ael(s, i), add = (Ael(e, i) - a0) / denom;
  Ael(t, i) = __dadd_rn(a0, __dmul_rn(f, add));  // no FMA: bit-exact with start + (k-s)*add
  a0 = Ael(s, i + n);
  add = (Ael(e, i + n) - a0) / denom;
  Ael(t, i + n) = __dadd_rn(a0, __dmul_rn(f, add));
  if (i < m) {
    double* Bb = Bm + (size_t)b * T * nx * m;
    double b0 = Bb[(size_t)s * nx * m + (size_t)i * nx + r];
    double addb = (Bb[(size_t)e * nx * m + (size_t)i * nx + r] - b0) / denom;
    Bb[(size_t)t * nx * m + (size_t)i * nx + r] = __dadd_rn(b0, __dmul_rn(f, addb));
  }
}

// iterative_error segment check (CheckDOFColumnError, KeyPointGenerator.cpp:550-640): one thread per segment
__global__ void k_iter_error(const double* __restrict__ A, const int* __restrict__ seg, int nseg, int T, int n, double thr,
                             int* __restrict__ good) {
  int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= nseg) return;
  int b = seg[4 * gid], dof = seg[4 * gid + 1], s = seg[4 * gid + 2], e = seg[4 * gid + 3];
  int mid = (s + e) / 2, nx = 2 * n;
  const double* Ab = A + (size_t)b * T * nx * nx;
  double err = 0.0;
  int counter = 0;
  for (int i = 0; i < 2; i++) {
    int c = i == 0 ? dof : dof + n;
    const double* As = Ab + (size_t)s * nx * nx + (size_t)c * nx;
    const double* Ae = Ab + (size_t)e * nx * nx + (size_t)c * nx;
    const double* Am = Ab + (size_t)mid * nx * nx + (size_t)c * nx;
    for (int j = n; j < nx; j++) {
      double approx = (As[j] + Ae[j]) / 2;
      double d = Am[j] - approx;
      counter++;
      err += __dmul_rn(d, d);
    }
  }
  double avg = counter > 0 ? err / counter : 0.0;
  good[gid] = avg < thr ? 1 : 0;
}

// GenerateKeyPointsIteratively, level-synchronous over all (b, dof) bins
static int keypoints_iterative(tokp_traj* tr) {
  int B = tr->B, T = tr->T, n = tr->n;
  struct Seg { int b, dof, s, e; };
  std::vector<Seg> list;
  for (int b = 0; b < B; b++)
    for (int i = 0; i < n; i++) list.push_back({b, i, 0, T - 1});
  std::vector<unsigned long long> computed((size_t)B * T, 0ull), need((size_t)B * T);
  // scratch of the bisection, allocated once per trajectory batch: need mask | row list | segments | verdicts
  size_t segcap = (size_t)B * n * (T + 1);
  if (!tr->ie_scratch) TOKP_CUDA(cudaMalloc(&tr->ie_scratch, (size_t)B * T * 8 + (size_t)B * T * 4 + segcap * 16 + segcap * 4));
  unsigned long long* d_need = static_cast<unsigned long long*>(tr->ie_scratch);
  int* d_rows = reinterpret_cast<int*>(d_need + (size_t)B * T);
  int* d_seg = d_rows + (size_t)B * T;
  int* d_good = d_seg + segcap * 4;
  int rc = TOKP_OK;
  while (!list.empty()) {
    // segments short enough are accepted before anything is computed (KeyPointGenerator.cpp:561-563)
    std::vector<Seg> check;
    for (auto& sg : list)
      if ((sg.e - sg.s) > tr->task.min_N) check.push_back(sg);
    if (check.empty()) break;
    std::fill(need.begin(), need.end(), 0ull);
    for (auto& sg : check) {
      int pts[3] = {sg.s, (sg.s + sg.e) / 2, sg.e};
      for (int p : pts) {
        size_t row = (size_t)sg.b * T + p;
        unsigned long long bit = 1ull << sg.dof;
        if (!(computed[row] & bit)) { need[row] |= bit; computed[row] |= bit; }
      }
    }
    std::vector<int> rows;
    for (size_t r = 0; r < need.size(); r++)
      if (need[r]) rows.push_back((int)r);
    if (!rows.empty()) {
      cudaMemcpyAsync(d_need, need.data(), need.size() * 8, cudaMemcpyHostToDevice, tr->stream);
      cudaMemcpyAsync(d_rows, rows.data(), rows.size() * 4, cudaMemcpyHostToDevice, tr->stream);
      rc = launch_fd(tr, d_need, d_rows, (int)rows.size(), 1, 1e-6);
      if (rc) break;
      for (size_t r = 0; r < need.size(); r++)
        if (need[r]) {
          int c = __builtin_popcountll(need[r]);
          int cm = __builtin_popcountll(need[r] & ((tr->m >= 64) ? ~0ull : ((1ull << tr->m) - 1)));
          tr->counters[0] += 1 + 2 * cm + 4 * c;
        }
    }
    std::vector<int> hseg(check.size() * 4), hgood(check.size());
    for (size_t i = 0; i < check.size(); i++) { hseg[4 * i] = check[i].b; hseg[4 * i + 1] = check[i].dof; hseg[4 * i + 2] = check[i].s; hseg[4 * i + 3] = check[i].e; }
    cudaMemcpyAsync(d_seg, hseg.data(), hseg.size() * 4, cudaMemcpyHostToDevice, tr->stream);
    k_iter_error<<<((int)check.size() + 127) / 128, 128, 0, tr->stream>>>(tr->A, d_seg, (int)check.size(), T, n, tr->task.iter_thr, d_good);
    g_launches++;
    cudaMemcpyAsync(hgood.data(), d_good, hgood.size() * 4, cudaMemcpyDeviceToHost, tr->stream);
    int hc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (!rows.empty()) cudaMemcpyAsync(hc, tr->fd_counts, sizeof(hc), cudaMemcpyDeviceToHost, tr->stream);
    if (cudaStreamSynchronize(tr->stream) != cudaSuccess) { set_error("iterative_error: CUDA failure"); rc = TOKP_ERR_CUDA; break; }
    if (!rows.empty() && tr->fd_last_lane_mode != 2) { adapt_fast_capacity(tr, hc[0] + hc[1], hc[2] + hc[3]); tr->counters[3] += hc[2] + hc[3]; }
    std::vector<Seg> next;
    for (size_t i = 0; i < check.size(); i++)
      if (!hgood[i]) {
        int mid = (check[i].s + check[i].e) / 2;
        next.push_back({check[i].b, check[i].dof, check[i].s, mid});
        next.push_back({check[i].b, check[i].dof, mid, check[i].e});
      }
    list.swap(next);
  }
  if (rc == TOKP_OK) {
    cudaMemcpyAsync(tr->kp_mask, computed.data(), computed.size() * 8, cudaMemcpyHostToDevice, tr->stream);
    cudaStreamSynchronize(tr->stream);
    long long cols = 0;  // Jacobians of this call: |C| / n per row, over every column the bisection differentiated
    for (auto mk : computed) cols += __builtin_popcountll(mk);
    tr->counters[1] = (int64_t)((double)cols / n * 1e6);
  }
  return rc;
}

__global__ void k_fill_int(int* p, int v, int count) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) p[i] = v;
}

int generate_keypoints(tokp_traj* tr) {
  int B = tr->B, T = tr->T, n = tr->n;
  const TaskDev& t = tr->task;
  int flags = 0;
  if (t.kp_method == TOKP_KP_SET_INTERVAL) {
    k_kp_set_interval<<<(B * T + 255) / 256, 256, 0, tr->stream>>>(tr->kp_mask, B, T, n, t.min_N);
    g_launches++;
  } else if (t.kp_method == TOKP_KP_ITERATIVE_ERROR) {
    tr->counters[0] = 0;
    int rc = keypoints_iterative(tr);
    if (rc) return rc;
  } else {
    cudaMemsetAsync(tr->kp_mask, 0, (size_t)B * T * 8, tr->stream);
    int mode = t.kp_method == TOKP_KP_ADAPTIVE_JERK ? 0 : (t.kp_method == TOKP_KP_ADAPTIVE_ACCEL ? 1 : 2);
    const double* thr = mode == 2 ? t.vel_thr : t.jerk_thr;
    cudaMemsetAsync(tr->kp_last, 0, (size_t)B * 8, tr->stream);
    k_kp_scan<<<(B * n + 63) / 64, 64, 0, tr->stream>>>(tr->kp_mask, tr->kp_last, tr->X, B, T, n, mode, t.min_N, t.max_N, thr, tr->dm.timestep);
    g_launches++;
    if (mode == 2) flags = 1;
  }
  k_fill_int<<<(B + 127) / 128, 128, 0, tr->stream>>>(tr->kp_flags, flags, B);  // no host buffer, no wait
  g_launches++;
  TOKP_CUDA(cudaGetLastError());
  return TOKP_OK;
}

int interpolate(tokp_traj* tr) {
  int B = tr->B, T = tr->T, n = tr->n;
  k_kp_links<<<(B * n + 63) / 64, 64, 0, tr->stream>>>(tr->kp_mask, tr->kp_prev, tr->kp_next, B, T, n);
  long long tot = (long long)B * T * n * tr->nx;
  k_interp<<<(int)((tot + 255) / 256), 256, 0, tr->stream>>>(tr->A, tr->Bm, tr->kp_mask, tr->kp_prev, tr->kp_next, B, T, n, tr->m);
  g_launches += 2;
  TOKP_CUDA(cudaGetLastError());
  return TOKP_OK;
}

}  // namespace tokp

using namespace tokp;

namespace tokp {
// host-side readback of the current key-point rows: lists in reference order (duplicates kept) + percentages
int read_keypoints(tokp_traj* tr, int* counts, int* indices, double* pct) {
  int B = tr->B, T = tr->T, n = tr->n;
  std::vector<unsigned long long> hm((size_t)B * T);
  std::vector<int> hf(B);
  std::vector<unsigned long long> hl(B);
  TOKP_CUDA(cudaMemcpyAsync(hl.data(), tr->kp_last, B * 8, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaMemcpyAsync(hm.data(), tr->kp_mask, hm.size() * 8, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaMemcpyAsync(hf.data(), tr->kp_flags, B * 4, cudaMemcpyDeviceToHost, tr->stream));
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  for (int b = 0; b < B; b++) {
    std::vector<int> cnt(n, 0);
    for (int t = 0; t < T; t++) {
      unsigned long long mk = hm[(size_t)b * T + t];
      if (t == T - 1 && (hf[b] & 1)) mk = hl[b];  // the scan's own hits; every dof is appended below
      int c = 0;
      int* row = indices ? indices + ((size_t)b * T + t) * 2 * n : nullptr;
      for (int j = 0; j < n; j++)
        if ((mk >> j) & 1ull) { if (row) row[c] = j; c++; cnt[j]++; }
      // velocity_change appends every dof again to the last row (KeyPointGenerator.cpp:724-727)
      if (t == T - 1 && (hf[b] & 1))
        for (int j = 0; j < n; j++) { if (row) row[c] = j; c++; cnt[j]++; }
      if (counts) counts[(size_t)b * T + t] = c;
    }
    if (pct)
      for (int j = 0; j < n; j++) pct[(size_t)b * n + j] = ((double)cnt[j] / (double)T) * 100;
  }
  return TOKP_OK;
}
}  // namespace tokp

extern "C" int tokp_keypoints(tokp_traj* tr, int* counts, int* indices, double* pct) {
  if (!tr) { set_error("tokp_keypoints: null trajectory"); return TOKP_ERR_ARG; }
  if (!tr->have_rollout && tr->task.kp_method != TOKP_KP_SET_INTERVAL) { set_error("tokp_keypoints: no nominal trajectory yet"); return TOKP_ERR_STATE; }
  int rc = generate_keypoints(tr);
  if (rc) return rc;
  if (!counts && !indices && !pct) return TOKP_OK;
  return read_keypoints(tr, counts, indices, pct);
}

extern "C" int tokp_traj_set_keypoints(tokp_traj* tr, int b, const int* counts, const int* indices) {
  if (!tr || !counts || !indices || b < 0 || b >= tr->B) { set_error("tokp_traj_set_keypoints: bad argument"); return TOKP_ERR_ARG; }
  int T = tr->T, n = tr->n;
  std::vector<unsigned long long> hm(T, 0ull);
  int p = 0;
  for (int t = 0; t < T; t++)
    for (int i = 0; i < counts[t]; i++) {
      int v = indices[p++];
      if (v < 0 || v >= n) { set_error("tokp_traj_set_keypoints: dof out of range"); return TOKP_ERR_ARG; }
      hm[t] |= 1ull << v;
    }
  TOKP_CUDA(cudaMemcpy(tr->kp_mask + (size_t)b * T, hm.data(), T * 8, cudaMemcpyHostToDevice));
  return TOKP_OK;
}

extern "C" int tokp_interpolate(tokp_traj* tr) {
  if (!tr) { set_error("tokp_interpolate: null trajectory"); return TOKP_ERR_ARG; }
  int rc = interpolate(tr);
  if (rc) return rc;
  TOKP_CUDA(cudaStreamSynchronize(tr->stream));
  return TOKP_OK;
}
