// ORACLE (test infrastructure only): extern "C" surface loaded by tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs via ctypes.  Nothing under trajoptkp_b200/
// may link or load this library.
#include <cstring>
#include <string>

#include "ok_ilqr.h"
#include "ok_init.h"

using namespace ok;

namespace {
thread_local std::string g_err;
template <class F>
int guard(F&& f) {
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return 1;
  }
}
template <class R>
int get_field(const Data<R>* d, const std::string& nm, double* out, int cap) {
  auto put = [&](const std::vector<R>& v, int cnt) {
    if (cnt > cap) return -cnt;
    for (int i = 0; i < cnt; i++) out[i] = to_double(v[i]);
    return cnt;
  };
  const Model* m = d->m;
  int nv = m->nv;
#define F(name, cnt) if (nm == #name) return put(d->name, (cnt));
  F(qpos, m->nq) F(qvel, nv) F(qacc, nv) F(qacc_warmstart, nv) F(ctrl, m->nu)
  F(xpos, 3 * m->nbody) F(xquat, 4 * m->nbody) F(xmat, 9 * m->nbody) F(xipos, 3 * m->nbody) F(ximat, 9 * m->nbody)
  F(xanchor, 3 * m->njnt) F(xaxis, 3 * m->njnt) F(geom_xpos, 3 * m->ngeom) F(geom_xmat, 9 * m->ngeom)
  F(subtree_com, 3 * m->nbody) F(cinert, 10 * m->nbody) F(cdof, 6 * nv) F(crb, 10 * m->nbody) F(qM, nv * nv)
  F(cvel, 6 * m->nbody) F(cdof_dot, 6 * nv) F(qfrc_passive, nv) F(qfrc_bias, nv) F(qfrc_actuator, nv)
  F(qfrc_smooth, nv) F(qacc_smooth, nv) F(qfrc_constraint, nv)
  F(efc_J, d->nefc * nv) F(efc_pos, d->nefc) F(efc_D, d->nefc) F(efc_R, d->nefc) F(efc_aref, d->nefc)
  F(efc_vel, d->nefc) F(efc_force, d->nefc) F(efc_diagApprox, d->nefc)
#undef F
  if (nm == "nefc") { out[0] = d->nefc; return 1; }
  if (nm == "ncon") { out[0] = d->ncon; return 1; }
  if (nm == "solver_niter") { out[0] = d->solver_niter; return 1; }
  if (nm == "warning") { out[0] = d->warning; return 1; }
  if (nm == "contact_dist") { for (int i = 0; i < d->ncon && i < cap; i++) out[i] = to_double(d->contact[i].dist); return d->ncon; }
  if (nm == "contact_pos") { for (int i = 0; i < d->ncon && 3 * i + 2 < cap; i++) for (int k = 0; k < 3; k++) out[3 * i + k] = to_double(d->contact[i].pos[k]); return 3 * d->ncon; }
  if (nm == "contact_frame") { for (int i = 0; i < d->ncon && 9 * i + 8 < cap; i++) for (int k = 0; k < 9; k++) out[9 * i + k] = to_double(d->contact[i].frame[k]); return 9 * d->ncon; }
  return -1000000;
}
}  // namespace

extern "C" {

const char* ok_last_error() { return g_err.c_str(); }

Model* ok_model_create(const uint8_t* blob, int64_t size) {
  Model* m = new Model();
  if (guard([&] { m->load(blob, size); })) { delete m; return nullptr; }
  return m;
}
void ok_model_destroy(Model* m) { delete m; }
int ok_snapshot_size(const Model* m) { return snapshot_size(m); }

Data<double>* ok_data_create(const Model* m) { return new Data<double>(m); }
void ok_data_destroy(Data<double>* d) { delete d; }
void ok_data_load_snapshot(Data<double>* d, const double* s) { load_snapshot(d->m, d, s); }
void ok_data_save_snapshot(const Data<double>* d, double* s) { save_snapshot(d->m, d, s); }
// iterations < 0: model options
void ok_data_step(Data<double>* d, int skipstage, int iterations, double tolerance) {
  SolverOpts opt = iterations < 0 ? model_opts(d->m) : SolverOpts{iterations, tolerance};
  step_skip(d->m, d, skipstage, opt);
}
void ok_data_forward(Data<double>* d, int skipstage, int iterations, double tolerance) {
  SolverOpts opt = iterations < 0 ? model_opts(d->m) : SolverOpts{iterations, tolerance};
  forward_skip(d->m, d, skipstage, opt);
}
int ok_data_get(const Data<double>* d, const char* name, double* out, int cap) { return get_field(d, name, out, cap); }

void ok_integrate_pos(const Model* m, double* qpos, const double* qvel, double dt) { integrate_pos(m, qpos, qvel, dt); }
void ok_differentiate_pos(const Model* m, double* qvel, double dt, const double* q1, const double* q2) {
  differentiate_pos(m, qvel, dt, q1, q2);
}

// flop count of one (possibly stage-skipping) step sequence: nominal full step then `skipstage` step
// from the same snapshot (what one perturbed world costs in the reference's schedule)
uint64_t ok_flops_step(const Model* m, const double* snapshot, int skipstage, int iterations, double tolerance) {
  Data<Counted> d(m);
  SolverOpts opt = iterations < 0 ? model_opts(m) : SolverOpts{iterations, tolerance};
  load_snapshot(m, &d, snapshot);
  step_skip(m, &d, STAGE_NONE, opt);
  load_snapshot(m, &d, snapshot);
  FlopCounter::n() = 0;
  step_skip(m, &d, skipstage, opt);
  return FlopCounter::n();
}

int ok_dynamics_derivs(const Model* m, const TaskDesc* td, const double* snapshot, const int* cols, int ncols, int central,
                       double eps, int iterations, double tolerance, double* A, double* B) {
  return guard([&] {
    Task task;
    task.init(m, td);
    Data<double> fd(m);
    SolverOpts opt = iterations < 0 ? model_opts(m) : SolverOpts{iterations, tolerance};
    dynamics_derivatives(m, task, &fd, snapshot, A, B, cols, ncols, central != 0, eps, opt);
  });
}

// algorithmic flops of one DynamicsDerivatives call (reference schedule incl. stage skipping)
uint64_t ok_flops_dynamics_derivs(const Model* m, const TaskDesc* td, const double* snapshot, const int* cols, int ncols,
                                  int iterations, double tolerance) {
  Task task;
  task.init(m, td);
  Data<Counted> fd(m);
  int nx = 2 * task.n;
  std::vector<double> A((size_t)nx * nx), B((size_t)nx * std::max(1, task.m));
  SolverOpts opt = iterations < 0 ? model_opts(m) : SolverOpts{iterations, tolerance};
  FlopCounter::n() = 0;
  dynamics_derivatives(m, task, &fd, snapshot, A.data(), B.data(), cols, ncols, true, 1e-6, opt);
  return FlopCounter::n();
}

int ok_residual_derivs(const Model* m, const TaskDesc* td, const double* snapshot, int central, double eps, double* rx,
                       double* ru) {
  return guard([&] {
    Task task;
    task.init(m, td);
    Data<double> fd(m);
    residual_derivatives(m, task, &fd, snapshot, rx, ru, central != 0, eps);
  });
}

int ok_residuals(const Model* m, const TaskDesc* td, const double* snapshot, double* r) {
  return guard([&] {
    Task task;
    task.init(m, td);
    Data<double> d(m);
    load_snapshot(m, &d, snapshot);
    residuals(task, &d, r);
  });
}

// key-points from a state trajectory X ((T+1) x 2n, row per time step).  Output: counts[T] and the
// concatenated index lists (capacity cap).  Returns total number of indices or -1.
int ok_keypoints(int method, const double* X, int n, int T, double dt, int min_N, int max_N, const double* jerk_thr,
                 const double* vel_thr, int* counts, int* indices, int cap) {
  KeyPoints kp;
  switch (method) {
    case KP_SET_INTERVAL: kp = keypoints_set_interval(n, T, min_N); break;
    case KP_ADAPTIVE_JERK: kp = keypoints_adaptive(jerk_profile(X, n, T, dt), n, T, min_N, max_N, jerk_thr); break;
    case KP_ADAPTIVE_ACCEL: kp = keypoints_adaptive(accel_profile(X, n, T), n, T, min_N, max_N, jerk_thr); break;
    case KP_VELOCITY_CHANGE: kp = keypoints_velocity_change(X, n, T, min_N, max_N, vel_thr); break;
    default: return -1;
  }
  int tot = 0;
  for (int t = 0; t < T; t++) {
    counts[t] = (int)kp[t].size();
    for (int v : kp[t]) {
      if (tot >= cap) return -1;
      indices[tot++] = v;
    }
  }
  return tot;
}

static KeyPoints unpack_kp(const int* counts, const int* indices, int T) {
  KeyPoints kp(T);
  int p = 0;
  for (int t = 0; t < T; t++)
    for (int i = 0; i < counts[t]; i++) kp[t].push_back(indices[p++]);
  return kp;
}

void ok_interpolate(const int* counts, const int* indices, int T, int n, int m, double* A, double* B) {
  interpolate_derivatives(unpack_kp(counts, indices, T), T, n, m, A, B);
}

void ok_percentages(const int* counts, const int* indices, int T, int n, double* pct) {
  auto p = keypoint_percentages(unpack_kp(counts, indices, T), n, T);
  for (int i = 0; i < n; i++) pct[i] = p[i];
}

// GenTestingData.cpp:177-220 for one BoxSweep / TwoDPushing trial (ok_init.h); returns warning bits or -1
int ok_init_sequence_ee(const Model* m, const TaskDesc* td, const double* start, const double* goal_xy, int setup_horizon,
                        int T, double* snapshot0, double* U, double* U_setup) {
  int warn = -1;
  if (guard([&] {
        Task task;
        task.init(m, td);
        warn = init_sequence_ee(m, task, start, goal_xy, setup_horizon, T, snapshot0, U, U_setup);
      }))
    return -1;
  return warn;
}

// ---- iLQR object ---------------------------------------------------------------------------
ILQR* ok_ilqr_create(const Model* m, const TaskDesc* td, int T, int threads) {
  ILQR* o = nullptr;
  if (guard([&] { o = new ILQR(m, td, T, threads); })) return nullptr;
  return o;
}
void ok_ilqr_destroy(ILQR* o) { delete o; }
void ok_ilqr_set_lambda(ILQR* o, double lambda) { o->lambda = lambda; }
void ok_ilqr_set_filter(ILQR* o, int method, double low_pass_a, const double* fir, int nfir) {
  o->filter_method = method;
  o->low_pass_a = low_pass_a;
  if (fir && nfir > 0) o->fir.assign(fir, fir + nfir);
}
void ok_ilqr_filter(ILQR* o) { o->filter_dynamics_matrices(); }
void ok_ilqr_set_A(ILQR* o, const double* A) { std::memcpy(o->A.data(), A, sizeof(double) * o->A.size()); }
void ok_ilqr_set_fd_model_opts(ILQR* o, int flag) { o->fd_use_model_opts = flag != 0; }
double ok_ilqr_rollout(ILQR* o, const double* snapshot0, const double* U) {
  o->old_cost = o->rollout_trajectory(snapshot0, U);
  o->new_cost = o->old_cost;
  return o->old_cost;
}
void ok_ilqr_generate_keypoints(ILQR* o) { o->generate_keypoints(); }
void ok_ilqr_dynamics_derivs(ILQR* o) { o->dynamics_derivatives_at_keypoints(); }
void ok_ilqr_generate_derivatives(ILQR* o) { o->generate_derivatives(); }
int ok_ilqr_backward(ILQR* o) { return o->backwards_pass() ? 1 : 0; }
double ok_ilqr_forward(ILQR* o, int thread_id, double alpha) { return o->forwards_pass(thread_id, alpha); }
void ok_ilqr_save_best(ILQR* o, int thread_id, double cost) {
  o->new_cost = cost;
  o->save_best_rollout(thread_id);
  o->update_nominal();
}
int ok_ilqr_iteration(ILQR* o, int* lambda_exit) {
  bool conv = false, le = false;
  o->iteration(conv, le);
  *lambda_exit = le;
  return conv;
}
void ok_ilqr_optimise(ILQR* o, const double* snapshot0, const double* U, int max_it, int min_it, double* U_out) {
  o->optimise(snapshot0, U, max_it, min_it, U_out);
}
// named array getter; returns count, or -count if cap too small, or -1000000 if unknown
int ok_ilqr_get(ILQR* o, const char* name, double* out, int cap) {
  std::string nm(name);
  auto put = [&](const std::vector<double>& v) {
    int c = (int)v.size();
    if (c > cap) return -c;
    std::memcpy(out, v.data(), sizeof(double) * c);
    return c;
  };
#define F(f) if (nm == #f) return put(o->f);
  F(saved) F(X_old) F(U_old) F(resid) F(A) F(B) F(r_x) F(r_u) F(l_x) F(l_xx) F(l_u) F(l_uu) F(K) F(k)
  F(last_percentages) F(cost_history) F(time_derivs_ms) F(time_bp_ms) F(time_fp_ms) F(pct_derivs) F(alphas_used)
#undef F
  if (nm == "scalars") {
    double v[] = {o->lambda, o->delta_J, o->old_cost, o->new_cost, o->initial_cost, o->cost_reduction,
                  (double)o->num_iterations, (double)o->fd_sims};
    int c = sizeof(v) / sizeof(double);
    if (c > cap) return -c;
    std::memcpy(out, v, sizeof(v));
    return c;
  }
  if (nm.rfind("rollout", 0) == 0) return put(o->rollout[nm[7] - '0']);
  return -1000000;
}
int ok_ilqr_get_keypoints(ILQR* o, int* counts, int* indices, int cap) {
  int tot = 0;
  for (int t = 0; t < o->T; t++) {
    counts[t] = (int)o->keypoints[t].size();
    for (int v : o->keypoints[t]) {
      if (tot >= cap) return -1;
      indices[tot++] = v;
    }
  }
  return tot;
}
void ok_ilqr_set_nominal(ILQR* o, const double* saved, const double* X, const double* U) {
  std::memcpy(o->saved.data(), saved, sizeof(double) * o->saved.size());
  std::memcpy(o->X_old.data(), X, sizeof(double) * o->X_old.size());
  std::memcpy(o->U_old.data(), U, sizeof(double) * o->U_old.size());
}

}  // extern "C"
