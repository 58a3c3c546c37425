// ORACLE (test infrastructure only).  Literal restatement of
//   Differentiator::DynamicsDerivatives  (src/Differentiator/Differentiator.cpp:8-462)
//   Differentiator::ResidualDerivatives  (src/Differentiator/Differentiator.cpp:464-663)
// on top of the oracle's MuJoCo-pipeline restatement.  Matrices are column-major with leading
// dimension = rows (Eigen default).  One `Data` per worker plays the role of fd_data[tid].
#pragma once
#include "ok_task.h"

namespace ok {

// FD solver options: MuJoCoHelper::InitModelForFiniteDifferencing (MuJoCoHelper.cpp:925-931)
inline SolverOpts fd_opts() { return SolverOpts{5, 0.0}; }
inline SolverOpts model_opts(const Model* m) { return SolverOpts{m->iterations, m->tolerance}; }

template <class R>
void get_full_state(const Model* m, const Data<R>* d, R* s) {  // mj_getState(mjSTATE_PHYSICS): qpos|qvel
  for (int i = 0; i < m->nq; i++) s[i] = d->qpos[i];
  for (int i = 0; i < m->nv; i++) s[m->nq + i] = d->qvel[i];
}

// A: nx x nx, B: nx x m (column-major); only columns {c, c+n} of A and {c | c<m} of B are written.
template <class R>
void dynamics_derivatives(const Model* m, const Task& task, Data<R>* fd, const double* snapshot, double* A, double* B,
                          const int* cols, int ncols, bool central_diff, double eps_d, const SolverOpts& opt,
                          long* count_integrations = nullptr) {
  int n = task.n, nu = task.m, nx = 2 * n, nq = m->nq, nv = m->nv;
  R eps(eps_d);
  std::vector<R> next_state(nx), next_plus(nx), next_minus(nx);
  std::vector<R> full(nq + nv, R(0.0)), full_pos(nq + nv, R(0.0)), full_minus(nq + nv, R(0.0));
  std::vector<R> dpos(nv), vel_diff(nv);
  std::vector<R> dqpos(nx * n, R(0.0)), dqvel(nx * n, R(0.0)), dctrl(nx * std::max(nu, 1), R(0.0));
  auto in_cols = [&](int i) {
    for (int c = 0; c < ncols; c++) if (cols[c] == i) return true;
    return false;
  };
  long nint = 0;
  // unperturbed full step; intermediates stay valid for the pre-step state
  load_snapshot(m, fd, snapshot);
  step_skip(m, fd, STAGE_NONE, opt);
  return_state_vector(task, fd, next_state.data());
  get_full_state(m, fd, full.data());
  load_snapshot(m, fd, snapshot);
  std::vector<R> u0(nu), v0(n);
  for (int i = 0; i < nu; i++) u0[i] = fd->ctrl[task.ctrl_adr[i]];
  for (int i = 0; i < n; i++) v0[i] = fd->qvel[task.dof_adr[i]];

  auto finish_column = [&](std::vector<R>& dst, int i, int mode) {
    // mode 0: central (minus -> plus, 2eps); 1: forward (nominal -> plus); 2: backward (minus -> nominal)
    const R* a = mode == 1 ? full.data() : full_minus.data();
    const R* b = mode == 2 ? full.data() : full_pos.data();
    R h = mode == 0 ? R(2.0) * eps : eps;
    differentiate_pos(m, vel_diff.data(), h, a, b);
    for (int j = 0; j < n; j++) dst[i * nx + j] = vel_diff[task.dof_adr[j]];
    for (int j = n; j < nx; j++) {
      if (mode == 0) dst[i * nx + j] = (next_plus[j] - next_minus[j]) / (R(2.0) * eps);
      else if (mode == 1) dst[i * nx + j] = (next_plus[j] - next_state[j]) / eps;
      else dst[i * nx + j] = (next_state[j] - next_minus[j]) / eps;
    }
  };

  // ---- controls (clamp-aware)
  for (int i = 0; i < nu; i++) {
    if (!in_cols(i)) continue;
    R up = u0[i] + eps;
    int nudge_forward = 1;
    if (up > R(task.ctrl_limits[2 * i + 1])) nudge_forward = 0;
    if (nudge_forward) {
      nint++;
      fd->ctrl[task.ctrl_adr[i]] = up;
      step_skip(m, fd, STAGE_VEL, opt);
      get_full_state(m, fd, full_pos.data());
      return_state_vector(task, fd, next_plus.data());
      load_snapshot(m, fd, snapshot);
    }
    R um = u0[i] - eps;
    int nudge_back;
    if (central_diff || !nudge_forward) {
      nudge_back = 1;
      if (um < R(task.ctrl_limits[2 * i])) nudge_back = 0;
    } else {
      nudge_back = 0;
    }
    if (nudge_back) {
      nint++;
      fd->ctrl[task.ctrl_adr[i]] = um;
      step_skip(m, fd, STAGE_VEL, opt);
      get_full_state(m, fd, full_minus.data());
      return_state_vector(task, fd, next_minus.data());
      load_snapshot(m, fd, snapshot);
    }
    if (nudge_forward && nudge_back) finish_column(dctrl, i, 0);
    else if (nudge_forward) finish_column(dctrl, i, 1);
    else if (nudge_back) finish_column(dctrl, i, 2);
  }
  // ---- velocities
  for (int i = 0; i < n; i++) {
    if (!in_cols(i)) continue;
    nint++;
    fd->qvel[task.dof_adr[i]] = v0[i] + eps;
    step_skip(m, fd, STAGE_POS, opt);
    get_full_state(m, fd, full_pos.data());
    return_state_vector(task, fd, next_plus.data());
    if (central_diff) {
      load_snapshot(m, fd, snapshot);
      fd->qvel[task.dof_adr[i]] = v0[i] - eps;
      step_skip(m, fd, STAGE_POS, opt);
      get_full_state(m, fd, full_minus.data());
      return_state_vector(task, fd, next_minus.data());
    }
    finish_column(dqvel, i, central_diff ? 0 : 1);
    load_snapshot(m, fd, snapshot);
  }
  // ---- positions
  for (int i = 0; i < n; i++) {
    if (!in_cols(i)) continue;
    nint++;
    zero(dpos.data(), nv);
    dpos[task.dof_adr[i]] = R(1.0);
    integrate_pos(m, fd->qpos.data(), dpos.data(), eps);
    step_skip(m, fd, STAGE_NONE, opt);
    get_full_state(m, fd, full_pos.data());
    return_state_vector(task, fd, next_plus.data());
    if (central_diff) {
      load_snapshot(m, fd, snapshot);
      integrate_pos(m, fd->qpos.data(), dpos.data(), -eps);
      step_skip(m, fd, STAGE_NONE, opt);
      get_full_state(m, fd, full_minus.data());
      return_state_vector(task, fd, next_minus.data());
    }
    finish_column(dqpos, i, central_diff ? 0 : 1);
    load_snapshot(m, fd, snapshot);
  }
  // ---- scatter the requested columns
  for (int c = 0; c < ncols; c++) {
    int col = cols[c];
    for (int j = 0; j < nx; j++) {
      A[col * nx + j] = to_double(dqpos[col * nx + j]);
      A[(col + n) * nx + j] = to_double(dqvel[col * nx + j]);
    }
    if (col < nu)
      for (int j = 0; j < nx; j++) B[col * nx + j] = to_double(dctrl[col * nx + j]);
  }
  if (count_integrations) *count_integrations = nint;
}

// r_x: num_resid x nx, r_u: num_resid x m, both column-major here ( r_x[j](i) == rx[i*num_resid + j] )
template <class R>
void residual_derivatives(const Model* m, const Task& task, Data<R>* fd, const double* snapshot, double* rx, double* ru,
                          bool central_diff, double eps_d) {
  int n = task.n, nu = task.m, nr = task.num_resid, nv = m->nv;
  R eps(eps_d);
  std::vector<R> r0(nr), rinc(nr), rdec(nr), dpos(nv);
  load_snapshot(m, fd, snapshot);
  residuals(task, fd, r0.data());
  std::vector<R> u0(nu), v0(n);
  for (int i = 0; i < nu; i++) u0[i] = fd->ctrl[task.ctrl_adr[i]];
  for (int i = 0; i < n; i++) v0[i] = fd->qvel[task.dof_adr[i]];
  for (int i = 0; i < nu; i++) {
    R up = u0[i] + eps;
    int nudge_forward = 1;
    if (up > R(task.ctrl_limits[2 * i + 1])) nudge_forward = 0;
    if (nudge_forward) {
      fd->ctrl[task.ctrl_adr[i]] = up;
      residuals(task, fd, rinc.data());
      load_snapshot(m, fd, snapshot);
    }
    R um = u0[i] - eps;
    int nudge_back;
    if (central_diff || !nudge_forward) {
      nudge_back = 1;
      if (um < R(task.ctrl_limits[2 * i])) nudge_back = 0;
    } else {
      nudge_back = 0;
    }
    if (nudge_back) {
      fd->ctrl[task.ctrl_adr[i]] = um;
      residuals(task, fd, rdec.data());
      load_snapshot(m, fd, snapshot);
    }
    for (int j = 0; j < nr; j++) {
      if (nudge_forward && nudge_back) ru[i * nr + j] = to_double((rinc[j] - rdec[j]) / (R(2.0) * eps));
      else if (nudge_forward) ru[i * nr + j] = to_double((rinc[j] - r0[j]) / eps);
      else if (nudge_back) ru[i * nr + j] = to_double((r0[j] - rdec[j]) / eps);
    }
  }
  for (int i = 0; i < n; i++) {
    fd->qvel[task.dof_adr[i]] = v0[i] + eps;
    residuals(task, fd, rinc.data());
    if (central_diff) {
      load_snapshot(m, fd, snapshot);
      fd->qvel[task.dof_adr[i]] = v0[i] - eps;
      residuals(task, fd, rdec.data());
    }
    for (int j = 0; j < nr; j++)
      rx[(i + n) * nr + j] = central_diff ? to_double((rinc[j] - rdec[j]) / (R(2.0) * eps)) : to_double((rinc[j] - r0[j]) / eps);
    load_snapshot(m, fd, snapshot);
  }
  for (int i = 0; i < n; i++) {
    zero(dpos.data(), nv);
    dpos[task.dof_adr[i]] = R(1.0);
    integrate_pos(m, fd->qpos.data(), dpos.data(), eps);
    residuals(task, fd, rinc.data());
    if (central_diff) {
      load_snapshot(m, fd, snapshot);
      integrate_pos(m, fd->qpos.data(), dpos.data(), -eps);
      residuals(task, fd, rdec.data());
    }
    for (int j = 0; j < nr; j++)
      rx[i * nr + j] = central_diff ? to_double((rinc[j] - rdec[j]) / (R(2.0) * eps)) : to_double((rinc[j] - r0[j]) / eps);
    load_snapshot(m, fd, snapshot);
  }
}

}  // namespace ok
