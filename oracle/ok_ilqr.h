// ORACLE (test infrastructure only).  Literal restatement (quirks included) of
//   Optimiser::GenerateDerivatives pipeline   src/Optimiser/Optimiser.cpp:80-338
//   Optimiser::SaveBestRollout / SaveSystemStateToRolloutData   src/Optimiser/Optimiser.cpp:416-468
//   iLQR::RolloutTrajectory :202-254, Optimise :269-410, Iteration :412-531,
//   BackwardsPassQuuRegularisation :535-634, UpdateLambda :636-657, CheckMatrixPD :659-670,
//   ForwardsPassParallel :824-934, UpdateNominal :936-948            (src/Optimiser/iLQR.cpp)
// Eigen's MatrixXd algebra is restated with plain loops; ldlt().solve(I) follows Eigen's pivoted
// LDLT (Eigen/src/Cholesky/LDLT.h, Eigen >= 3.3 - not vendored in /root/reference).
#pragma once
#include <atomic>
#include <chrono>
#include <cmath>
#include <thread>

#include "ok_deriv.h"
#include "ok_keypoints.h"

namespace ok {

// Eigen LLT success test (CheckMatrixPD): fails iff a pivot is <= 0
inline bool check_matrix_pd(const double* Ain, int n) {
  std::vector<double> L(Ain, Ain + n * n);  // column-major, lower used
  for (int k = 0; k < n; k++) {
    double x = L[k * n + k];
    for (int j = 0; j < k; j++) x -= L[j * n + k] * L[j * n + k];
    if (x <= 0.0) return false;
    x = std::sqrt(x);
    L[k * n + k] = x;
    for (int i = k + 1; i < n; i++) {
      double s = L[k * n + i];
      for (int j = 0; j < k; j++) s -= L[j * n + i] * L[j * n + k];
      L[k * n + i] = s / x;
    }
  }
  return true;
}

// Eigen LDLT (lower, diagonal pivoting) followed by solve(I).  A, Ainv column-major n x n.
inline void ldlt_inverse(const double* Ain, int n, double* Ainv) {
  std::vector<double> mat(Ain, Ain + n * n);
  auto M = [&](int r, int c) -> double& { return mat[c * n + r]; };
  std::vector<int> transp(n);
  bool zero_matrix = false;
  for (int k = 0; k < n; k++) {
    int big = k;
    double bigv = std::fabs(M(k, k));
    for (int i = k + 1; i < n; i++)
      if (std::fabs(M(i, i)) > bigv) { bigv = std::fabs(M(i, i)); big = i; }
    transp[k] = big;
    if (k != big) {
      int s = n - big - 1;
      for (int c = 0; c < k; c++) std::swap(M(k, c), M(big, c));
      for (int r = 0; r < s; r++) std::swap(M(big + 1 + r, k), M(big + 1 + r, big));
      std::swap(M(k, k), M(big, big));
      for (int i = k + 1; i < big; i++) std::swap(M(i, k), M(big, i));
    }
    int rs = n - k - 1;
    if (k > 0) {
      std::vector<double> temp(k);
      for (int j = 0; j < k; j++) temp[j] = M(j, j) * M(k, j);
      double s = 0;
      for (int j = 0; j < k; j++) s += M(k, j) * temp[j];
      M(k, k) -= s;
      for (int i = 0; i < rs; i++) {
        double t = 0;
        for (int j = 0; j < k; j++) t += M(k + 1 + i, j) * temp[j];
        M(k + 1 + i, k) -= t;
      }
    }
    double akk = M(k, k);
    bool valid = std::fabs(akk) > 0.0;
    if (k == 0 && !valid) {
      for (int j = 0; j < n; j++) transp[j] = j;
      zero_matrix = true;
      break;
    }
    if (rs > 0 && valid)
      for (int i = 0; i < rs; i++) M(k + 1 + i, k) /= akk;
  }
  const double tol = 2.2250738585072014e-308;  // numeric_limits<double>::min()
  for (int c = 0; c < n; c++) {
    std::vector<double> x(n, 0.0);
    x[c] = 1.0;
    if (!zero_matrix) {
      for (int k = 0; k < n; k++) std::swap(x[k], x[transp[k]]);
      for (int i = 0; i < n; i++)
        for (int j = 0; j < i; j++) x[i] -= M(i, j) * x[j];
      for (int i = 0; i < n; i++) x[i] = std::fabs(M(i, i)) > tol ? x[i] / M(i, i) : 0.0;
      for (int i = n - 1; i >= 0; i--)
        for (int j = i + 1; j < n; j++) x[i] -= M(j, i) * x[j];
      for (int k = n - 1; k >= 0; k--) std::swap(x[k], x[transp[k]]);
    } else {
      std::fill(x.begin(), x.end(), 0.0);
    }
    for (int r = 0; r < n; r++) Ainv[c * n + r] = x[r];
  }
}

struct ILQR {
  const Model* mdl;
  Task task;
  int n, m, nx, nr, T, S;
  int num_threads;
  static const int kRollouts = 6;
  bool fd_use_model_opts = false;  // debug: run FD with the model's solver options

  // nominal trajectory = saved_systems_state_list (T+1 snapshots)
  std::vector<double> saved;      // (T+1) x S
  std::vector<double> X_old;      // (T+1) x nx
  std::vector<double> U_old;      // T x m
  std::vector<double> resid;      // (T+1) x nr
  std::vector<double> A, B;       // T x nx*nx, T x nx*m
  std::vector<double> r_x, r_u;   // (T+1) x nr*nx, (T+1) x nr*m
  std::vector<double> l_x, l_xx, l_u, l_uu;  // (T+1) x ...
  std::vector<double> K, k;       // T x m*nx, T x m
  std::vector<double> rollout[kRollouts];  // (T+1) x S each, zero-initialised (iLQR::Resize :150-186)
  KeyPoints keypoints;
  std::vector<double> last_percentages;
  std::vector<std::unique_ptr<Data<double>>> fd;  // fd_data[tid]
  std::unique_ptr<Data<double>> main_data;

  double lambda = 0.1, max_lambda = 10.0, min_lambda = 0.0001, lambda_factor = 10.0;
  double eps_converge = 0.02;
  double delta_J = 0, old_cost = 0, new_cost = 0, initial_cost = 0, cost_reduction = 0;
  bool cost_reduced_last_iter = true;
  int num_iterations = 0;
  std::vector<double> cost_history, time_derivs_ms, time_bp_ms, time_fp_ms, pct_derivs, alphas_used;
  long fd_sims = 0;

  ILQR(const Model* model, const TaskDesc* td, int horizon, int threads) : mdl(model) {
    task.init(model, td);
    n = task.n; m = task.m; nx = 2 * n; nr = task.num_resid; T = horizon; S = snapshot_size(model);
    num_threads = threads > 0 ? threads : std::max(1, (int)std::thread::hardware_concurrency() - 1);
    saved.assign((size_t)(T + 1) * S, 0.0);
    X_old.assign((size_t)(T + 1) * nx, 0.0);
    U_old.assign((size_t)T * m, 0.0);
    resid.assign((size_t)(T + 1) * nr, 0.0);
    A.assign((size_t)T * nx * nx, 0.0); B.assign((size_t)T * nx * m, 0.0);
    r_x.assign((size_t)(T + 1) * nr * nx, 0.0); r_u.assign((size_t)(T + 1) * nr * m, 0.0);
    l_x.assign((size_t)(T + 1) * nx, 0.0); l_xx.assign((size_t)(T + 1) * nx * nx, 0.0);
    l_u.assign((size_t)(T + 1) * m, 0.0); l_uu.assign((size_t)(T + 1) * m * m, 0.0);
    K.assign((size_t)T * m * nx, 0.0); k.assign((size_t)T * m, 0.0);
    for (auto& r : rollout) r.assign((size_t)(T + 1) * S, 0.0);
    int nfd = std::max(num_threads, kRollouts);
    for (int i = 0; i < nfd; i++) fd.emplace_back(new Data<double>(model));
    main_data.reset(new Data<double>(model));
    last_percentages.assign(n, 0.0);
  }

  double* snap(int t) { return &saved[(size_t)t * S]; }
  void state_of(const Data<double>* d, double* x) { return_state_vector(task, d, x); }
  void set_controls(Data<double>* d, const double* u) { for (int i = 0; i < m; i++) d->ctrl[task.ctrl_adr[i]] = u[i]; }

  // iLQR::RolloutTrajectory(d, save_states=true, U)
  double rollout_trajectory(const double* snapshot0, const double* U) {
    double cost = 0.0;
    Data<double>* d = main_data.get();
    load_snapshot(mdl, d, snapshot0);
    state_of(d, &X_old[0]);
    save_snapshot(mdl, d, snap(0));
    SolverOpts opt = model_opts(mdl);
    for (int i = 0; i < T; i++) {
      set_controls(d, U + (size_t)i * m);
      step_skip(mdl, d, STAGE_NONE, opt);
      residuals(task, d, &resid[(size_t)i * nr]);
      cost += cost_function(task, &resid[(size_t)i * nr], i == T - 1);
      state_of(d, &X_old[(size_t)(i + 1) * nx]);
      for (int c = 0; c < m; c++) U_old[(size_t)i * m + c] = d->ctrl[task.ctrl_adr[c]];
      save_snapshot(mdl, d, snap(i + 1));
    }
    initial_cost = cost;
    cost_history.push_back(cost);
    return cost;
  }

  // KeypointGenerator::GenerateKeyPoints
  void generate_keypoints() {
    double dt = mdl->timestep;
    switch (task.kp_method) {
      case KP_SET_INTERVAL: keypoints = keypoints_set_interval(n, T, task.min_N); break;
      case KP_ADAPTIVE_JERK:
        keypoints = keypoints_adaptive(jerk_profile(X_old.data(), n, T, dt), n, T, task.min_N, task.max_N, task.jerk_thr.data());
        break;
      case KP_ADAPTIVE_ACCEL:
        keypoints = keypoints_adaptive(accel_profile(X_old.data(), n, T), n, T, task.min_N, task.max_N, task.jerk_thr.data());
        break;
      case KP_VELOCITY_CHANGE:
        keypoints = keypoints_velocity_change(X_old.data(), n, T, task.min_N, task.max_N, task.vel_thr.data());
        break;
      case KP_ITERATIVE_ERROR: {
        SolverOpts opt = fd_use_model_opts ? model_opts(mdl) : fd_opts();
        keypoints = keypoints_iterative(n, T, task.min_N, task.iter_thr, A.data(), [&](int t, int dof) {
          long ni = 0;
          dynamics_derivatives(mdl, task, fd[0].get(), snap(t), &A[(size_t)t * nx * nx], &B[(size_t)t * nx * m], &dof, 1, true, 1e-6, opt, &ni);
          fd_sims += ni + 1;
        });
        break;
      }
    }
    last_percentages = keypoint_percentages(keypoints, n, T);
  }

  // Optimiser::ComputeDynamicsDerivativesAtKeypoints + WorkerComputeDerivatives
  void dynamics_derivatives_at_keypoints() {
    std::vector<int> time_indices;
    for (int t = 0; t < (int)keypoints.size(); t++)
      if (!keypoints[t].empty()) time_indices.push_back(t);
    std::atomic<int> next(0);
    std::atomic<long> sims(0);
    SolverOpts opt = fd_use_model_opts ? model_opts(mdl) : fd_opts();
    auto worker = [&](int tid) {
      for (;;) {
        int it = next.fetch_add(1);
        if (it >= (int)time_indices.size()) break;
        int t = time_indices[it];
        long ni = 0;
        dynamics_derivatives(mdl, task, fd[tid].get(), snap(t), &A[(size_t)t * nx * nx], &B[(size_t)t * nx * m],
                             keypoints[t].data(), (int)keypoints[t].size(), true, 1e-6, opt, &ni);
        sims += ni + 1;
      }
    };
    run_threads(worker);
    fd_sims += sims.load();
  }

  template <class F>
  void run_threads(F&& worker) {
    if (num_threads <= 1) { worker(0); return; }
    std::vector<std::thread> pool;
    for (int i = 0; i < num_threads; i++) pool.emplace_back(worker, i);
    for (auto& th : pool) th.join();
  }

  // Optimiser::ComputeResidualDerivatives (t = 0..T)
  void residual_derivatives_all() {
    std::atomic<int> next(0);
    auto worker = [&](int tid) {
      for (;;) {
        int t = next.fetch_add(1);
        if (t >= T + 1) break;
        residual_derivatives(mdl, task, fd[tid].get(), snap(t), &r_x[(size_t)t * nr * nx], &r_u[(size_t)t * nr * m], true, 1e-6);
      }
    };
    run_threads(worker);
  }

  // MT::CostDerivativesFromResiduals
  void cost_derivs_from_residuals(int t, bool terminal) {
    double* lx = &l_x[(size_t)t * nx];
    double* lxx = &l_xx[(size_t)t * nx * nx];
    double* lu = &l_u[(size_t)t * m];
    double* luu = &l_uu[(size_t)t * m * m];
    std::fill(lx, lx + nx, 0.0); std::fill(lxx, lxx + nx * nx, 0.0);
    std::fill(lu, lu + m, 0.0); std::fill(luu, luu + m * m, 0.0);
    const double* r = &resid[(size_t)t * nr];
    const double* rx = &r_x[(size_t)t * nr * nx];
    const double* ru = &r_u[(size_t)t * nr * m];
    for (int i = 0; i < nr; i++) {
      double w = terminal ? task.weight_terminal[i] : task.weight[i];
      for (int a = 0; a < nx; a++) lx[a] += w * 2 * r[i] * rx[a * nr + i];
      for (int b = 0; b < nx; b++)
        for (int a = 0; a < nx; a++) lxx[b * nx + a] += w * 2 * rx[a * nr + i] * rx[b * nr + i];
      for (int a = 0; a < m; a++) lu[a] += w * 2 * r[i] * ru[a * nr + i];
      for (int b = 0; b < m; b++)
        for (int a = 0; a < m; a++) luu[b * m + a] += 2 * w * ru[a * nr + i] * ru[b * nr + i];
    }
  }

  // Optimiser::FilterDynamicsMatrices (Optimiser.cpp:340-372): every entry of the velocity rows of A is treated as a
  // signal over t = 0..T-1 and replaced by its filtered version.  low_pass (Optimiser.cpp:374-390):
  // y_t = (1-a) y_{t-1} + a (x_t + x_{t-1})/2 with y_{-1} = x_{-1} = x_0.  FIR (Optimiser.cpp:392-406): causal
  // convolution, taps that would reach before t = 0 are dropped.
  int filter_method = 0;  // 0 none, 1 low_pass, 2 FIR
  double low_pass_a = 0.25;                               // include/Optimiser/Optimiser.h:219
  std::vector<double> fir = {0.1, 0.15, 0.5, 0.15, 0.1};  // include/Optimiser/Optimiser.h:220
  void filter_dynamics_matrices() {
    std::vector<double> x(T), y(T);
    for (int row = n; row < nx; row++)
      for (int col = 0; col < nx; col++) {
        for (int t = 0; t < T; t++) x[t] = A[(size_t)t * nx * nx + (size_t)col * nx + row];
        if (filter_method == 1) {
          double yprev = x[0], xprev = x[0];
          for (int t = 0; t < T; t++) {
            y[t] = ((1 - low_pass_a) * yprev) + low_pass_a * ((x[t] + xprev) / 2);
            xprev = x[t];
            yprev = y[t];
          }
        } else {
          for (int t = 0; t < T; t++) {
            double acc = 0;
            for (int j = 0; j < (int)fir.size() && j <= t; j++) acc += x[t - j] * fir[j];
            y[t] = acc;
          }
        }
        for (int t = 0; t < T; t++) A[(size_t)t * nx * nx + (size_t)col * nx + row] = y[t];
      }
  }

  // Optimiser::GenerateDerivatives
  void generate_derivatives() {
    generate_keypoints();
    if (task.kp_method != KP_ITERATIVE_ERROR) dynamics_derivatives_at_keypoints();
    interpolate_derivatives(keypoints, T, n, m, A.data(), B.data());
    residual_derivatives_all();
    for (int t = 0; t < T; t++) cost_derivs_from_residuals(t, false);
    cost_derivs_from_residuals(T - 1, true);
    double avg = 0.0;
    for (int i = 0; i < n; i++) avg += last_percentages[i];
    pct_derivs.push_back(avg / n);
    if (filter_method != 0) filter_dynamics_matrices();
  }

  // iLQR::BackwardsPassQuuRegularisation
  bool backwards_pass() {
    std::vector<double> Vx(&l_x[(size_t)(T - 1) * nx], &l_x[(size_t)(T - 1) * nx] + nx);
    std::vector<double> Vxx(&l_xx[(size_t)(T - 1) * nx * nx], &l_xx[(size_t)(T - 1) * nx * nx] + nx * nx);
    std::vector<double> Qx(nx), Qu(m), Qxx(nx * nx), Quu(m * m), Qux(m * nx), Quu_reg(m * m), Quu_inv(m * m);
    std::vector<double> VA(nx * nx), VB(nx * m), tmp_mx(m * nx), tmp_m(m);
    int pd_counter = 0;
    delta_J = 0.0;
    for (int t = T - 1; t >= 0; t--) {
      pd_counter++;
      const double* At = &A[(size_t)t * nx * nx];
      const double* Bt = &B[(size_t)t * nx * m];
      // Q_x = l_x + A' V_x ; Q_u = l_u + B' V_x
      for (int i = 0; i < nx; i++) {
        double s = 0;
        for (int r = 0; r < nx; r++) s += At[i * nx + r] * Vx[r];
        Qx[i] = l_x[(size_t)t * nx + i] + s;
      }
      for (int i = 0; i < m; i++) {
        double s = 0;
        for (int r = 0; r < nx; r++) s += Bt[i * nx + r] * Vx[r];
        Qu[i] = l_u[(size_t)t * m + i] + s;
      }
      // VA = V_xx A ; VB = V_xx B
      for (int c = 0; c < nx; c++)
        for (int r = 0; r < nx; r++) {
          double s = 0;
          for (int q = 0; q < nx; q++) s += Vxx[q * nx + r] * At[c * nx + q];
          VA[c * nx + r] = s;
        }
      for (int c = 0; c < m; c++)
        for (int r = 0; r < nx; r++) {
          double s = 0;
          for (int q = 0; q < nx; q++) s += Vxx[q * nx + r] * Bt[c * nx + q];
          VB[c * nx + r] = s;
        }
      // (A' V_xx) A evaluated as A' (V_xx A): differs from Eigen's left-to-right order only in rounding
      for (int c = 0; c < nx; c++)
        for (int r = 0; r < nx; r++) {
          double s = 0;
          for (int q = 0; q < nx; q++) s += At[r * nx + q] * VA[c * nx + q];
          Qxx[c * nx + r] = l_xx[(size_t)t * nx * nx + c * nx + r] + s;
        }
      for (int c = 0; c < m; c++)
        for (int r = 0; r < m; r++) {
          double s = 0;
          for (int q = 0; q < nx; q++) s += Bt[r * nx + q] * VB[c * nx + q];
          Quu[c * m + r] = l_uu[(size_t)t * m * m + c * m + r] + s;
        }
      for (int c = 0; c < nx; c++)
        for (int r = 0; r < m; r++) {
          double s = 0;
          for (int q = 0; q < nx; q++) s += Bt[r * nx + q] * VA[c * nx + q];
          Qux[c * m + r] = s;
        }
      Quu_reg = Quu;
      for (int i = 0; i < m; i++) Quu_reg[i * m + i] += lambda;
      if (pd_counter >= 100) {
        if (!check_matrix_pd(Quu_reg.data(), m)) return false;
        pd_counter = 0;
      }
      ldlt_inverse(Quu_reg.data(), m, Quu_inv.data());
      double* kt = &k[(size_t)t * m];
      double* Kt = &K[(size_t)t * m * nx];
      for (int r = 0; r < m; r++) {
        double s = 0;
        for (int q = 0; q < m; q++) s += Quu_inv[q * m + r] * Qu[q];
        kt[r] = -s;
      }
      for (int c = 0; c < nx; c++)
        for (int r = 0; r < m; r++) {
          double s = 0;
          for (int q = 0; q < m; q++) s += Quu_inv[q * m + r] * Qux[c * m + q];
          Kt[c * m + r] = -s;
        }
      // V_x = Q_x + K'(Q_uu k) + K' Q_u + Q_ux' k
      for (int r = 0; r < m; r++) {
        double s = 0;
        for (int q = 0; q < m; q++) s += Quu[q * m + r] * kt[q];
        tmp_m[r] = s;
      }
      for (int i = 0; i < nx; i++) {
        double a = 0, b = 0, c = 0;
        for (int r = 0; r < m; r++) { a += Kt[i * m + r] * tmp_m[r]; b += Kt[i * m + r] * Qu[r]; c += Qux[i * m + r] * kt[r]; }
        Vx[i] = Qx[i] + a + b + c;
      }
      // V_xx = Q_xx + K'(Q_uu K) + K' Q_ux + Q_ux' K
      for (int c = 0; c < nx; c++)
        for (int r = 0; r < m; r++) {
          double s = 0;
          for (int q = 0; q < m; q++) s += Quu[q * m + r] * Kt[c * m + q];
          tmp_mx[c * m + r] = s;
        }
      for (int c = 0; c < nx; c++)
        for (int r = 0; r < nx; r++) {
          double a = 0, b = 0, d = 0;
          for (int q = 0; q < m; q++) { a += Kt[r * m + q] * tmp_mx[c * m + q]; b += Kt[r * m + q] * Qux[c * m + q]; d += Qux[r * m + q] * Kt[c * m + q]; }
          Vxx[c * nx + r] = Qxx[c * nx + r] + a + b + d;
        }
      for (int c = 0; c < nx; c++)
        for (int r = 0; r < c; r++) {
          double s = (Vxx[c * nx + r] + Vxx[r * nx + c]) / 2;
          Vxx[c * nx + r] = s; Vxx[r * nx + c] = s;
        }
      double d1 = 0, d2 = 0;
      for (int r = 0; r < m; r++) { d1 += kt[r] * Qu[r]; d2 += kt[r] * tmp_m[r]; }
      delta_J += d1;
      delta_J += d2;
    }
    return true;
  }

  bool update_lambda(bool valid) {
    bool lambda_exit = false;
    if (!valid) lambda *= lambda_factor; else lambda /= lambda_factor;
    if (lambda > max_lambda) { lambda = max_lambda; lambda_exit = true; }
    if (lambda < min_lambda) lambda = min_lambda;
    return lambda_exit;
  }

  // iLQR::ForwardsPassParallel (n == dof_quat branch)
  double forwards_pass(int thread_id, double alpha) {
    double cost = 0.0;
    Data<double>* d = fd[thread_id].get();
    load_snapshot(mdl, d, snap(0));
    SolverOpts opt = model_opts(mdl);
    std::vector<double> X_new(nx), fb(nx), U_new(m), r(nr);
    for (int t = 0; t < T; t++) {
      state_of(d, X_new.data());
      for (int i = 0; i < nx; i++) fb[i] = X_new[i] - X_old[(size_t)t * nx + i];
      const double* Kt = &K[(size_t)t * m * nx];
      for (int i = 0; i < m; i++) {
        double g = 0;
        for (int c = 0; c < nx; c++) g += Kt[c * m + i] * fb[c];
        U_new[i] = U_old[(size_t)t * m + i] + (alpha * k[(size_t)t * m + i]) + g;
        if (U_new[i] > task.ctrl_limits[2 * i + 1]) U_new[i] = task.ctrl_limits[2 * i + 1];
        if (U_new[i] < task.ctrl_limits[2 * i]) U_new[i] = task.ctrl_limits[2 * i];
      }
      set_controls(d, U_new.data());
      residuals(task, d, r.data());
      cost += cost_function(task, r.data(), t == T - 1);
      step_skip(mdl, d, STAGE_NONE, opt);
      if (t != 0) save_snapshot(mdl, d, &rollout[thread_id][(size_t)t * S]);
    }
    return cost;
  }

  // Optimiser::SaveBestRollout (with its one-step index shift)
  void save_best_rollout(int thread_id) {
    int nq = mdl->nq, nv = mdl->nv, nu = mdl->nu, nb = mdl->nbody;
    int off_ctrl = 1 + nq + 4 * nv + 6 * nb;
    Data<double>* d = main_data.get();
    for (int t = 0; t < T; t++) {
      double* dst = snap(t + 1);
      const double* src = &rollout[thread_id][(size_t)(t + 1) * S];
      for (int i = 0; i < off_ctrl; i++) dst[i] = src[i];
      double* dctrl = snap(t) + off_ctrl;
      const double* sctrl = &rollout[thread_id][(size_t)t * S] + off_ctrl;
      for (int i = 0; i < nu; i++) dctrl[i] = sctrl[i];
      // residuals[t+1] from saved[t+1] as it stands now (its ctrl is overwritten on the next loop turn)
      load_snapshot(mdl, d, snap(t + 1));
      residuals(task, d, &resid[(size_t)(t + 1) * nr]);
    }
  }

  void update_nominal() {
    Data<double>* d = main_data.get();
    int off_ctrl = 1 + mdl->nq + 4 * mdl->nv + 6 * mdl->nbody;
    for (int t = 0; t < T; t++) {
      load_snapshot(mdl, d, snap(t + 1));
      state_of(d, &X_old[(size_t)(t + 1) * nx]);
      const double* c = snap(t) + off_ctrl;
      for (int i = 0; i < m; i++) U_old[(size_t)t * m + i] = c[task.ctrl_adr[i]];
    }
    old_cost = new_cost;
  }

  static double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
  }

  // iLQR::Iteration
  void iteration(bool& converged, bool& lambda_exit) {
    double t0 = now_ms();
    if (cost_reduced_last_iter) generate_derivatives();
    time_derivs_ms.push_back(now_ms() - t0);
    t0 = now_ms();
    bool valid = false;
    lambda_exit = false;
    while (!valid) {
      valid = backwards_pass();
      lambda_exit = update_lambda(valid);
      if (lambda_exit) break;
    }
    time_bp_ms.push_back(now_ms() - t0);
    if (lambda_exit) return;
    t0 = now_ms();
    double alphas[kRollouts], results[kRollouts];
    for (int i = 1; i < kRollouts + 1; i++) {
      double lin = (double)i / kRollouts;
      alphas[i - 1] = lin * lin;
    }
    {
      std::vector<std::thread> pool;
      for (int i = 0; i < kRollouts; i++) pool.emplace_back([&, i]() { results[i] = forwards_pass(i, alphas[i]); });
      for (auto& th : pool) th.join();
    }
    int best = 0;
    for (int i = 1; i < kRollouts; i++)
      if (results[i] < results[best]) best = i;
    alphas_used.push_back(alphas[best]);
    if (results[best] < old_cost) {
      new_cost = results[best];
      save_best_rollout(best);
    } else {
      new_cost = old_cost;
    }
    time_fp_ms.push_back(now_ms() - t0);
    converged = ((old_cost - new_cost) / new_cost) < eps_converge;
    if (new_cost < old_cost) {
      update_nominal();
      cost_reduced_last_iter = true;
    } else {
      cost_reduced_last_iter = false;
      lambda *= lambda_factor;
      lambda *= lambda_factor;
      if (lambda > max_lambda) lambda = max_lambda;
    }
    cost_history.push_back(new_cost);
  }

  // iLQR::Optimise; U_out receives U_old (T x m)
  void optimise(const double* snapshot0, const double* U_init, int max_iterations, int min_iterations, double* U_out) {
    cost_history.clear(); time_derivs_ms.clear(); time_bp_ms.clear(); time_fp_ms.clear(); pct_derivs.clear();
    alphas_used.clear();
    num_iterations = 0;
    old_cost = rollout_trajectory(snapshot0, U_init);
    new_cost = old_cost;
    initial_cost = old_cost;
    cost_reduced_last_iter = true;
    for (int i = 0; i < max_iterations; i++) {
      num_iterations++;
      bool lambda_exit = false, converged = false;
      iteration(converged, lambda_exit);
      if (converged && (i >= min_iterations)) break;
      if (lambda_exit) break;
    }
    cost_reduction = 1 - (new_cost / initial_cost);
    if (U_out) std::copy(U_old.begin(), U_old.end(), U_out);
  }
};

}  // namespace ok
