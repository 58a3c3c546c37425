// ORACLE (test infrastructure only).  Literal restatement of the reference's key-point rules and
// derivative interpolation (src/KeyPointGenerator/KeyPointGenerator.cpp:319-339 set_interval,
// :341-382 adaptive, :449-640 iterative_error, :642-728 velocity_change, :730-808 profiles,
// :814-838 percentages, :840-954 InterpolateDerivatives).  X[t] is the (dof_quat + n) state; all
// configs in scope have dof_quat == n so X[t] = [pos(n); vel(n)].
#pragma once
#include <algorithm>
#include <cmath>
#include <functional>
#include <vector>

namespace ok {

typedef std::vector<std::vector<int>> KeyPoints;

inline KeyPoints keypoints_set_interval(int n, int T, int min_N) {
  KeyPoints kp;
  std::vector<int> full(n), empty;
  for (int i = 0; i < n; i++) full[i] = i;
  for (int t = 0; t < T - 1; t++) kp.push_back(t % min_N == 0 ? full : empty);
  kp.push_back(full);
  return kp;
}

// GenerateJerkProfile: profile[t][j], rows T-2 and T-1 zero
inline std::vector<double> jerk_profile(const double* X, int n, int T, double dt) {
  int nx = 2 * n;
  std::vector<double> prof((size_t)T * n, 0.0);
  for (int t = 0; t < T - 2; t++) {
    const double *s1 = X + (size_t)t * nx, *s2 = X + (size_t)(t + 1) * nx, *s3 = X + (size_t)(t + 2) * nx;
    for (int j = 0; j < n; j++) {
      double a1 = (s2[j + n] - s1[j + n]) / dt;
      double a2 = (s3[j + n] - s2[j + n]) / dt;
      prof[(size_t)t * n + j] = std::fabs((a2 - a1) / dt);
    }
  }
  return prof;
}

// GenerateAccellerationProfile: T-1 rows of raw velocity differences
inline std::vector<double> accel_profile(const double* X, int n, int T) {
  int nx = 2 * n;
  std::vector<double> prof((size_t)T * n, 0.0);
  for (int t = 0; t < T - 1; t++)
    for (int j = 0; j < n; j++) prof[(size_t)t * n + j] = X[(size_t)(t + 1) * nx + j + n] - X[(size_t)t * nx + j + n];
  return prof;
}

inline KeyPoints keypoints_adaptive(const std::vector<double>& prof, int n, int T, int min_N, int max_N, const double* thr) {
  KeyPoints kp;
  std::vector<int> full(n);
  for (int i = 0; i < n; i++) full[i] = i;
  kp.push_back(full);
  std::vector<int> last(n, 0);
  for (int t = 1; t < T - 1; t++) {
    std::vector<int> row;
    for (int j = 0; j < n; j++) {
      if ((t - last[j]) >= min_N) {
        if (prof[(size_t)t * n + j] > thr[j]) { row.push_back(j); last[j] = t; }
      }
      if ((t - last[j]) >= max_N) { row.push_back(j); last[j] = t; }
    }
    kp.push_back(row);
  }
  kp.push_back(full);
  return kp;
}

inline KeyPoints keypoints_velocity_change(const double* X, int n, int T, int min_N, int max_N, const double* thr) {
  int nx = 2 * n;
  auto vel = [&](int t, int i) { return X[(size_t)t * nx + n + i]; };
  KeyPoints kp;
  std::vector<int> full(n);
  for (int i = 0; i < n; i++) full[i] = i;
  kp.push_back(full);
  std::vector<int> counter(n, 0);
  std::vector<double> last_val(n, 0.0), last_dir(n, 0.0);
  for (int t = 1; t < T; t++) {
    std::vector<int> row;
    for (int i = 0; i < n; i++) {
      counter[i]++;
      double dir = vel(t, i) - vel(t - 1, i);
      last_val[i] += std::fabs(vel(t, i));
      if (counter[i] >= min_N) {
        if (std::fabs(last_val[i]) > thr[i]) { row.push_back(i); last_val[i] = 0.0; counter[i] = 0; continue; }
      }
      if (counter[i] >= min_N) {
        if (dir * last_dir[i] < 0) { row.push_back(i); last_val[i] = 0.0; counter[i] = 0; continue; }
      } else {
        last_dir[i] = dir;
      }
      if (counter[i] >= max_N) { row.push_back(i); last_val[i] = 0.0; counter[i] = 0; continue; }
    }
    kp.push_back(row);
  }
  for (int i = 0; i < n; i++) kp[T - 1].push_back(i);
  return kp;
}

// GenerateKeyPointsIteratively; `compute_column(t, dof)` must fill A[t](:, dof), A[t](:, dof+n), B[t](:, dof)
inline KeyPoints keypoints_iterative(int n, int T, int min_N, double threshold, double* A,
                                     const std::function<void(int, int)>& compute_column) {
  int nx = 2 * n;
  std::vector<std::vector<char>> computed(n, std::vector<char>(T, 0));
  auto check = [&](int start, int end, int dof) -> bool {
    int mid = (start + end) / 2;
    if ((end - start) <= min_N) return true;
    if (!computed[dof][start]) { compute_column(start, dof); computed[dof][start] = 1; }
    if (!computed[dof][mid]) { compute_column(mid, dof); computed[dof][mid] = 1; }
    if (!computed[dof][end]) { compute_column(end, dof); computed[dof][end] = 1; }
    double error_sum = 0.0;
    int counter = 0;
    int colidx[2] = {dof, dof + n};
    for (int i = 0; i < 2; i++) {
      for (int j = n; j < nx; j++) {
        const double* As = A + (size_t)start * nx * nx + (size_t)colidx[i] * nx;
        const double* Ae = A + (size_t)end * nx * nx + (size_t)colidx[i] * nx;
        const double* Am = A + (size_t)mid * nx * nx + (size_t)colidx[i] * nx;
        double approx = (As[j] + Ae[j]) / 2;
        double diff = Am[j] - approx;
        counter++;
        error_sum += diff * diff;
      }
    }
    double average_error = counter > 0 ? error_sum / counter : 0.0;
    return average_error < threshold;
  };
  for (int i = 0; i < n; i++) {
    std::vector<std::pair<int, int>> list{{0, T - 1}}, sub;
    bool complete = false;
    while (!complete) {
      bool all_ok = true;
      for (auto& seg : list) {
        int mid = (seg.first + seg.second) / 2;
        if (!check(seg.first, seg.second, i)) {
          all_ok = false;
          sub.push_back({seg.first, mid});
          sub.push_back({mid, seg.second});
        }
      }
      if (all_ok) complete = true;
      list = sub;
      sub.clear();
    }
  }
  KeyPoints kp(T);
  for (int t = 0; t < T; t++)
    for (int j = 0; j < n; j++)
      if (computed[j][t]) kp[t].push_back(j);
  return kp;
}

// ComputePercentageDerivatives
inline std::vector<double> keypoint_percentages(const KeyPoints& kp, int n, int T) {
  std::vector<int> cnt(n, 0);
  for (int t = 0; t < T; t++)
    for (int v : kp[t])
      if (v >= 0 && v < n) cnt[v]++;
  std::vector<double> pct(n);
  for (int i = 0; i < n; i++) pct[i] = ((double)cnt[i] / (double)T) * 100;
  return pct;
}

// InterpolateDerivatives (A: T x [nx x nx], B: T x [nx x m], column-major blocks)
inline void interpolate_derivatives(const KeyPoints& kp, int T, int n, int m, double* A, double* B) {
  int nx = 2 * n;
  std::vector<int> start(n, 0);
  auto Acol = [&](int t, int c) { return A + (size_t)t * nx * nx + (size_t)c * nx; };
  auto Bcol = [&](int t, int c) { return B + (size_t)t * nx * m + (size_t)c * nx; };
  std::vector<double> add1(nx), add2(nx), addB(nx);
  for (int t = 1; t < T; t++) {
    for (int i = 0; i < n; i++) {
      const std::vector<int>& columns = kp[t];
      if (columns.empty()) continue;
      for (size_t j = 0; j < columns.size(); j++) {
        if (i != columns[j]) continue;
        int s = start[i];
        double denom = (double)(t - s);
        for (int r = 0; r < nx; r++) {
          add1[r] = (Acol(t, i)[r] - Acol(s, i)[r]) / denom;
          add2[r] = (Acol(t, i + n)[r] - Acol(s, i + n)[r]) / denom;
          if (i < m) addB[r] = (Bcol(t, i)[r] - Bcol(s, i)[r]) / denom;
        }
        for (int k = s + 1; k < t; k++) {
          double f = (double)(k - s);
          for (int r = 0; r < nx; r++) {
            Acol(k, i)[r] = Acol(s, i)[r] + (f * add1[r]);
            Acol(k, i + n)[r] = Acol(s, i + n)[r] + (f * add2[r]);
            if (i < m) Bcol(k, i)[r] = Bcol(s, i)[r] + (f * addB[r]);
          }
        }
        start[i] = t;
      }
    }
  }
}

}  // namespace ok
