// Third part of the host mirror: Optimiser base storage and iLQR (include/Optimiser/Optimiser.h:49-242,
// include/Optimiser/iLQR.h; src/Optimiser/iLQR.cpp:202-410).  One tokp_optimise call per Optimise(): the whole
// iteration loop (key-points, FD derivatives, interpolation, cost derivatives, backward pass, 6-alpha line search,
// lambda schedule, convergence test) runs behind the C ABI on the device.
#pragma once
#include "tokp_shim_classes.h"

class Optimiser {
 public:
  Optimiser(std::shared_ptr<ModelTranslator> mt, std::shared_ptr<MuJoCoHelper> helper, std::shared_ptr<Differentiator> diff)
      : activeModelTranslator(mt), MuJoCo_helper(helper), differentiator(diff) {
    dof = mt->dof; num_ctrl = mt->num_ctrl;
  }
  virtual ~Optimiser() = default;
  virtual double RolloutTrajectory(mjData* d, bool save_states, std::vector<MatrixXd> initial_controls) = 0;
  virtual std::vector<MatrixXd> Optimise(mjData* d, std::vector<MatrixXd> initial_controls, int max_iterations,
                                         int min_iterations, int horizon_length) = 0;
  // Optimiser::SetCurrentKeypointMethod (Optimiser.cpp:47-50)
  void SetCurrentKeypointMethod(keypoint_method m) { keypoint_generator->SetKeypointMethod(m); }
  // Optimiser::setFIRFilter (Optimiser.cpp:408-414)
  void setFIRFilter(std::vector<double> _FIRCoefficients) { FIRCoefficients = _FIRCoefficients; }
  // Optimiser::FilterDynamicsMatrices (Optimiser.cpp:340-372) on the public A: the velocity rows are filtered along t
  // on the device; an unknown filteringMethod is reported as in the reference and leaves A untouched
  void FilterDynamicsMatrices() {
    tokp_traj* h = differentiator->traj->h;
    if (PushFilter() == TOKP_FILTER_NONE) { std::cerr << "Filtering method not recognised" << std::endl; return; }
    int nx = 2 * dof, T = (int)A.size();
    std::vector<double> a((size_t)T * nx * nx);
    for (int t = 0; t < T; t++) std::memcpy(&a[(size_t)t * nx * nx], A[t].data(), sizeof(double) * nx * nx);
    tokp_shim::check(tokp_traj_set(h, 0, "A", a.data(), (int)a.size()), "set A");
    tokp_shim::check(tokp_filter_dynamics(h), "FilterDynamicsMatrices");
    if (tokp_traj_get(h, 0, "A", a.data(), (int)a.size()) != (int)a.size()) throw std::runtime_error("FilterDynamicsMatrices: readback of A failed");
    for (int t = 0; t < T; t++) std::memcpy(A[t].data(), &a[(size_t)t * nx * nx], sizeof(double) * nx * nx);
  }
  // filteringMethod / lowPassACoefficient / FIRCoefficients -> device handle; returns the TOKP_FILTER_* code in force
  int PushFilter() {
    int code = filteringMethod == "low_pass" ? TOKP_FILTER_LOW_PASS : (filteringMethod == "FIR" ? TOKP_FILTER_FIR : TOKP_FILTER_NONE);
    tokp_shim::check(tokp_traj_set_filter(differentiator->traj->h, code, lowPassACoefficient, FIRCoefficients.data(),
                                          (int)FIRCoefficients.size()), "set_filter");
    return code;
  }

  // public storage, names as in include/Optimiser/Optimiser.h:166-242
  std::vector<MatrixXd> A, B, X_old, U_old, l_x, l_xx, l_u, l_uu;
  std::vector<double> cost_history, time_get_derivs_ms, time_backwards_pass_ms, time_forwardsPass_ms;
  double initial_cost = 0, cost_reduction = 0, opt_time_ms = 0, avg_percent_derivs = 0, avg_dofs = 0;
  double lambda = 0.1, new_cost = 0, old_cost = 0;
  int num_iterations = 0, dof = 0, num_ctrl = 0, horizon_length = 0;
  bool verbose_output = false;
  std::string filteringMethod = "none";                               // include/Optimiser/Optimiser.h:173
  double lowPassACoefficient = 0.25;                                  // :219
  std::vector<double> FIRCoefficients = {0.1, 0.15, 0.5, 0.15, 0.1};  // :220
  std::shared_ptr<KeypointGenerator> keypoint_generator;
  std::shared_ptr<ModelTranslator> activeModelTranslator;
  std::shared_ptr<MuJoCoHelper> MuJoCo_helper;
  std::shared_ptr<Differentiator> differentiator;
};

class iLQR : public Optimiser {
 public:
  iLQR(std::shared_ptr<ModelTranslator> mt, std::shared_ptr<MuJoCoHelper> helper, std::shared_ptr<Differentiator> diff,
       int horizon)
      : Optimiser(mt, helper, diff) {
    Resize(mt->dof, mt->num_ctrl, horizon);
    keypoint_generator = std::make_shared<KeypointGenerator>(diff, helper, dof, horizon);
    keypoint_generator->SetKeypointMethod(mt->current_kp);
  }
  // iLQR::Resize (iLQR.cpp:37-200): host mirrors only; the device arena was sized by tokp_traj_create
  void Resize(int new_dofs, int new_ctrl, int new_horizon) {
    dof = new_dofs; num_ctrl = new_ctrl; horizon_length = new_horizon;
    int nx = 2 * dof, T = new_horizon;
    if (T != differentiator->traj->T) throw std::runtime_error("iLQR::Resize: horizon differs from the device arena");
    A.assign(T, MatrixXd(nx, nx)); B.assign(T, MatrixXd(nx, num_ctrl));
    X_old.assign(T + 1, MatrixXd(nx, 1)); U_old.assign(T, MatrixXd(num_ctrl, 1));
    l_x.assign(T + 1, MatrixXd(nx, 1)); l_xx.assign(T + 1, MatrixXd(nx, nx));
    l_u.assign(T + 1, MatrixXd(num_ctrl, 1)); l_uu.assign(T + 1, MatrixXd(num_ctrl, num_ctrl));
  }
  // iLQR::RolloutTrajectory (iLQR.cpp:202-254)
  double RolloutTrajectory(mjData* d, bool /*save_states*/, std::vector<MatrixXd> initial_controls) override {
    std::vector<double> U = Flatten(initial_controls);
    double cost = 0;
    tokp_shim::check(tokp_rollout(differentiator->traj->h, d->v.data(), U.data(), &cost), "RolloutTrajectory");
    PullTrajectory();
    return cost;
  }
  // Optimiser::GenerateDerivatives (Optimiser.cpp:80-108)
  void GenerateDerivatives() {
    PushFilter();
    tokp_shim::check(tokp_generate_derivatives(differentiator->traj->h), "GenerateDerivatives");
    keypoint_generator->PullAB(A, B);
  }
  // iLQR::Optimise (iLQR.cpp:269-410); controls are returned by value as in the reference
  std::vector<MatrixXd> Optimise(mjData* d, std::vector<MatrixXd> initial_controls, int max_iterations, int min_iterations,
                                 int horizon) override {
    if (horizon != horizon_length) throw std::runtime_error("iLQR::Optimise: horizon differs from Resize()");
    std::vector<double> U = Flatten(initial_controls), Uo((size_t)horizon * num_ctrl);
    double stats[8];
    PushFilter();
    tokp_shim::check(tokp_optimise(differentiator->traj->h, d->v.data(), U.data(), max_iterations, min_iterations, &lambda,
                                   Uo.data(), stats), "Optimise");
    initial_cost = stats[0]; new_cost = stats[1]; num_iterations = (int)stats[2]; lambda = stats[3];
    avg_percent_derivs = stats[4]; avg_dofs = dof;
    time_get_derivs_ms = {stats[5]}; time_backwards_pass_ms = {stats[6]}; time_forwardsPass_ms = {stats[7]};
    opt_time_ms = stats[5] + stats[6] + stats[7];
    cost_reduction = 1 - (new_cost / initial_cost);  // iLQR.cpp:344
    PullTrajectory();
    std::vector<MatrixXd> out(horizon, MatrixXd(num_ctrl, 1));
    for (int t = 0; t < horizon; t++)
      for (int i = 0; i < num_ctrl; i++) out[t](i, 0) = Uo[(size_t)t * num_ctrl + i];
    return out;
  }

 private:
  std::vector<double> Flatten(const std::vector<MatrixXd>& c) const {
    if ((int)c.size() < horizon_length) throw std::runtime_error("control sequence shorter than the horizon");
    std::vector<double> U((size_t)horizon_length * num_ctrl);
    for (int t = 0; t < horizon_length; t++)
      for (int i = 0; i < num_ctrl; i++) U[(size_t)t * num_ctrl + i] = c[t](i, 0);
    return U;
  }
  void PullTrajectory() {
    tokp_traj* h = differentiator->traj->h;
    int nx = 2 * dof, T = horizon_length;
    std::vector<double> X((size_t)(T + 1) * nx), U((size_t)T * (num_ctrl ? num_ctrl : 1));
    tokp_traj_get(h, 0, "X_old", X.data(), (int)X.size());
    if (num_ctrl) tokp_traj_get(h, 0, "U_old", U.data(), (int)U.size());
    for (int t = 0; t <= T; t++) std::memcpy(X_old[t].data(), &X[(size_t)t * nx], sizeof(double) * nx);
    for (int t = 0; t < T && num_ctrl; t++) std::memcpy(U_old[t].data(), &U[(size_t)t * num_ctrl], sizeof(double) * num_ctrl);
    // saved_systems_state_list[t] <- device snapshots (UpdateNominal reads them, iLQR.cpp:936-948)
    auto& list = MuJoCo_helper->saved_systems_state_list;
    while ((int)list.size() < T + 1) list.emplace_back(new mjData{std::vector<double>(MuJoCo_helper->S, 0.0)});
    for (int t = 0; t <= T; t++) tokp_traj_get_snapshot(h, 0, t, list[t]->v.data());
  }
};
