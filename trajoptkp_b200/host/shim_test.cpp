// Driver for the C++ host mirror: runs the reference's call sequence (main.cpp:100-165 wiring, then
// RolloutTrajectory -> GenerateKeyPoints -> DynamicsDerivatives -> InterpolateDerivatives -> Optimise) on the acrobot
// task through the shim classes and prints one JSON object that tests/test_gpu_shim.py compares with the ctypes path
// and the oracle.  usage: shim_test <acrobot.xml> [--no-gpu]
//                        shim_test <boxSweep.xml> --boxsweep <input.txt>   (a target config: 7 arm joints + the goal
//                        body's x, y, z, vector targets, aux bodies, iterative_error key-points; the input file holds
//                        T, the start snapshot, the goal and the initial controls - written by tests/test_gpu_shim.py)
#include <cmath>
#include <cstdio>
#include <fstream>

#include "tokp_shim_optimiser.h"

static void print_vec(const char* name, const std::vector<double>& v, bool last = false) {
  std::printf("\"%s\": [", name);
  for (size_t i = 0; i < v.size(); i++) std::printf("%s%.17g", i ? ", " : "", v[i]);
  std::printf("]%s\n", last ? "" : ",");
}

// BASELINE configs[2] through the shim: InitModelTranslator with rigid bodies -> RolloutTrajectory -> SetCurrentKeypointMethod
// (iterative_error) -> GenerateKeyPoints -> GenerateDerivatives -> DynamicsDerivatives (per-call) -> Optimise
static int run_boxsweep(const char* xml, const char* input) {
  std::ifstream in(input);
  if (!in) { std::fprintf(stderr, "shim_test: cannot read %s\n", input); return 2; }
  int T = 0;
  in >> T;
  auto helper = std::make_shared<MuJoCoHelper>();
  helper->InitSimulator(0.008, xml);  // box_sweep.yaml: timeStep 0.008
  mjData start = *helper->master_reset_data;
  for (int i = 0; i < helper->S; i++) in >> start.v[i];
  double gx, gy;
  in >> gx >> gy;
  std::vector<MatrixXd> U0(T, MatrixXd(7, 1));
  for (int t = 0; t < T; t++) for (int i = 0; i < 7; i++) in >> U0[t](i, 0);
  if (!in) { std::fprintf(stderr, "shim_test: short input file\n"); return 2; }
  // TaskConfigs/rigid_body_manipulation/box_sweep.yaml
  keypoint_method kp;
  kp.name = "iterative_error"; kp.min_N = 1; kp.max_N = 50; kp.iterative_error_threshold = 1e-5;
  kp.jerk_thresholds.assign(7, 1.0);
  kp.velocity_change_thresholds.assign(7, 0.5);
  rigid_body goal;
  goal.name = "goal";
  for (int k = 0; k < 3; k++) { goal.active_linear_dof[k] = true; goal.linear_jerk_threshold[k] = 1.0; goal.linear_vel_change_threshold[k] = 0.05; }
  std::vector<std::string> joints, acts;
  for (int i = 1; i <= 7; i++) { joints.push_back("panda0_joint" + std::to_string(i)); acts.push_back("panda0_joint" + std::to_string(i)); }
  auto mt = std::make_shared<ModelTranslator>();
  mt->InitModelTranslator(helper, TOKP_TASK_BOXSWEEP, joints, acts, {goal}, {gx, gy, 0.0, 0.0, 0.05}, {0.0, 0.2, 0.0},
                          {1000.0, 100.0, 0.0}, kp, {"goal", "franka_gripper"});
  std::printf("{\"dof\": %d, \"num_ctrl\": %d, \"num_resid\": %d,\n", mt->dof, mt->num_ctrl, (int)mt->resid_weight.size());
  {
    MatrixXd x = mt->ReturnStateVector(&start), r;
    std::vector<double> xs(x.data(), x.data() + 2 * mt->dof);
    print_vec("start_state", xs);
    mt->Residuals(&start, r);
    print_vec("start_residuals", {r(0, 0), r(1, 0), r(2, 0)});
  }
  auto traj = std::make_shared<TrajHandle>(mt, T);
  auto diff = std::make_shared<Differentiator>(mt, helper, traj);
  iLQR opt(mt, helper, diff, T);
  double cost = opt.RolloutTrajectory(&start, true, U0);
  std::printf("\"rollout_cost\": %.17g,\n", cost);
  opt.SetCurrentKeypointMethod(mt->current_kp);  // thresholds of all 10 state DoFs (robot joints, then the body's)
  opt.keypoint_generator->GenerateKeyPoints(opt.X_old, opt.A, opt.B);
  std::vector<double> rows;
  int tq = -1;
  for (int t = 0; t < T; t++) {
    rows.push_back((double)opt.keypoint_generator->keypoints[t].size());
    if (tq < 0 && t > 0 && (int)opt.keypoint_generator->keypoints[t].size() == mt->dof) tq = t;
  }
  if (tq < 0) tq = T - 1;
  print_vec("keypoint_counts", rows);
  print_vec("percentages", opt.keypoint_generator->last_percentages);
  opt.GenerateDerivatives();
  const int nx = 2 * mt->dof;
  MatrixXd At(nx, nx), Bt(nx, 7);
  std::vector<int> cols;
  for (int i = 0; i < mt->dof; i++) cols.push_back(i);
  diff->DynamicsDerivatives(At, Bt, cols, tq, 0, true, 1e-6);
  double dmax = 0;
  for (int i = 0; i < nx; i++) {
    for (int j = 0; j < nx; j++) dmax = std::fmax(dmax, std::fabs(At(i, j) - opt.A[tq](i, j)));
    for (int j = 0; j < 7; j++) dmax = std::fmax(dmax, std::fabs(Bt(i, j) - opt.B[tq](i, j)));
  }
  std::printf("\"t_query\": %d, \"cols_vs_batched_maxabs\": %.3g,\n", tq, dmax);
  print_vec("A_t", std::vector<double>(At.data(), At.data() + nx * nx));
  opt.lambda = 0.1;
  auto Uopt = opt.Optimise(&start, U0, 2, 1, T);
  std::vector<double> uo;
  for (auto& u : Uopt) for (int i = 0; i < 7; i++) uo.push_back(u(i, 0));
  print_vec("U_opt", uo);
  std::printf("\"initial_cost\": %.17g, \"final_cost\": %.17g, \"iterations\": %d}\n", opt.initial_cost, opt.new_cost,
              opt.num_iterations);
  return 0;
}

int main(int argc, char** argv) {
  if (argc < 2) { std::fprintf(stderr, "usage: shim_test model.xml [--no-gpu | --boxsweep input.txt]\n"); return 2; }
  if (argc > 3 && std::string(argv[2]) == "--boxsweep") {
    try {
      return run_boxsweep(argv[1], argv[3]);
    } catch (const std::exception& e) {
      std::fprintf(stderr, "shim_test: %s\n", e.what());
      return 1;
    }
  }
  bool no_gpu = argc > 2 && std::string(argv[2]) == "--no-gpu";
  const int T = 50;
  try {
    auto helper = std::make_shared<MuJoCoHelper>();
    helper->InitSimulator(0.01, argv[1]);
    keypoint_method kp;
    kp.name = "set_interval"; kp.min_N = 5; kp.max_N = 50;
    kp.jerk_thresholds = {0.1, 0.1}; kp.velocity_change_thresholds = {6.0, 6.0};
    auto mt = std::make_shared<ModelTranslator>();
    // TaskConfigs/toys/acrobot.yaml: joints shoulder, elbow; actuator shoulder; residual weights / terminal weights
    mt->InitModelTranslator(helper, TOKP_TASK_ACROBOT, {"shoulder", "elbow"}, {"shoulder"},
                            {0.0, 0.0, 0.0, 0.0, 0.0}, {0.0, 0.0, 0.001, 0.001, 100.0}, {100.0, 100.0, 1.0, 1.0, 100.0}, kp);
    std::printf("{\"dof\": %d, \"num_ctrl\": %d, \"nq\": %d, \"S\": %d,\n", mt->dof, mt->num_ctrl, helper->nq, helper->S);
    {
      // ModelTranslator gather / scatter and host residuals (no device needed)
      mjData d = *helper->master_reset_data;
      MatrixXd x(4, 1);
      x(0, 0) = 0.5; x(1, 0) = -0.25; x(2, 0) = 1.5; x(3, 0) = -2.0;
      MatrixXd u(1, 1);
      u(0, 0) = 0.75;
      bool ok = mt->SetStateVector(x, &d) && mt->SetControlVector(u, &d);
      MatrixXd xr = mt->ReturnStateVector(&d), pr = mt->ReturnPositionVector(&d), vr = mt->ReturnVelocityVector(&d);
      MatrixXd lim = mt->ReturnControlLimits();
      MatrixXd r;
      mt->Residuals(&d, r);
      MatrixXd bad(3, 1);
      bool rejected = !mt->SetStateVector(bad, &d);
      std::printf("\"scatter_ok\": %d, \"size_mismatch_rejected\": %d,\n", ok ? 1 : 0, rejected ? 1 : 0);
      print_vec("state_roundtrip", {xr(0, 0), xr(1, 0), xr(2, 0), xr(3, 0), pr(1, 0), vr(1, 0), mt->ReturnControlVector(&d)(0, 0)});
      print_vec("ctrl_limits", {lim(0, 0), lim(1, 0)});
      print_vec("residuals", {r(0, 0), r(1, 0), r(2, 0), r(3, 0), r(4, 0)});
      std::printf("\"cost_running\": %.17g, \"cost_terminal\": %.17g,\n", mt->CostFunction(r, false), mt->CostFunction(r, true));
    }
    if (no_gpu) {
      // without a device the arena cannot be created: the error must surface, nothing falls back to the host
      try {
        TrajHandle th(mt, T);
        std::printf("\"traj_create\": \"unexpected success\"}\n");
        return 1;
      } catch (const std::exception& e) {
        std::printf("\"traj_create_error\": \"%s\"}\n", "raised");
        return 0;
      }
    }
    auto traj = std::make_shared<TrajHandle>(mt, T);
    auto diff = std::make_shared<Differentiator>(mt, helper, traj);
    iLQR opt(mt, helper, diff, T);
    mjData start = *helper->master_reset_data;
    helper->qpos(&start)[0] = 0.3;
    helper->qpos(&start)[1] = -0.2;
    std::vector<MatrixXd> U0(T, MatrixXd(1, 1));
    for (int t = 0; t < T; t++) U0[t](0, 0) = 0.2 * std::sin(0.3 * t);
    double cost = opt.RolloutTrajectory(&start, true, U0);
    std::printf("\"rollout_cost\": %.17g,\n", cost);
    opt.keypoint_generator->GenerateKeyPoints(opt.X_old, opt.A, opt.B);
    std::vector<double> rows;
    for (int t = 0; t < T; t++) rows.push_back((double)opt.keypoint_generator->keypoints[t].size());
    print_vec("keypoint_counts", rows);
    print_vec("percentages", opt.keypoint_generator->last_percentages);
    opt.GenerateDerivatives();
    // per-call drop-in at key-point t = 5 must reproduce the batched result
    MatrixXd A5(4, 4), B5(4, 1);
    diff->DynamicsDerivatives(A5, B5, {0, 1}, 5, 0, true, 1e-6);
    double dmax = 0;
    for (int i = 0; i < 4; i++) {
      for (int j = 0; j < 4; j++) dmax = std::fmax(dmax, std::fabs(A5(i, j) - opt.A[5](i, j)));
      dmax = std::fmax(dmax, std::fabs(B5(i, 0) - opt.B[5](i, 0)));
    }
    std::printf("\"cols_vs_batched_maxabs\": %.3g,\n", dmax);
    std::vector<double> a5(A5.data(), A5.data() + 16);
    print_vec("A5", a5);
    // Keypoints_Test.cpp:272-289: A[1] == A[0] + (A[3]-A[0])/3 after interpolation with key-points {0, 3, ...}
    std::vector<std::vector<int>> kps(T);
    for (int t = 0; t < T; t++) if (t % 3 == 0 || t == T - 1) kps[t] = {0, 1};
    std::vector<MatrixXd> A2 = opt.A, B2 = opt.B;
    std::vector<std::vector<MatrixXd>> rx, ru;
    opt.keypoint_generator->InterpolateDerivatives(kps, T, A2, B2, rx, ru, false, 1);
    int exact = 1;
    for (int i = 0; i < 4; i++)
      for (int j = 0; j < 4; j++) {
        double add = (opt.A[3](i, j) - opt.A[0](i, j)) / 3;
        if (A2[1](i, j) != opt.A[0](i, j) + 1 * add || A2[2](i, j) != opt.A[0](i, j) + 2 * add) exact = 0;
      }
    std::printf("\"interpolation_identity_exact\": %d,\n", exact);
    // Differentiator::ResidualDerivatives drop-in at t = 5 (r_x[j] nx x 1, r_u[j] m x 1)
    std::vector<MatrixXd> r_x, r_u;
    diff->ResidualDerivatives(r_x, r_u, 5, 0, true, 1e-6);
    std::vector<double> rxf, ruf;
    for (auto& r : r_x) for (int c = 0; c < 4; c++) rxf.push_back(r(c, 0));
    for (auto& r : r_u) ruf.push_back(r(0, 0));
    print_vec("r_x5", rxf);
    print_vec("r_u5", ruf);
    // Optimiser::FilterDynamicsMatrices on a copy of A (low_pass, then FIR with the default taps)
    {
      std::vector<MatrixXd> keep = opt.A;
      opt.filteringMethod = "low_pass";
      opt.FilterDynamicsMatrices();
      std::vector<double> lp;
      for (int t = 0; t < T; t++) lp.push_back(opt.A[t](2, 1));
      print_vec("A_lowpass_21", lp);
      int untouched = 1;
      for (int t = 0; t < T; t++) for (int j = 0; j < 4; j++) if (opt.A[t](0, j) != keep[t](0, j) || opt.A[t](1, j) != keep[t](1, j)) untouched = 0;
      std::printf("\"filter_position_rows_untouched\": %d,\n", untouched);
      opt.A = keep;
      opt.filteringMethod = "FIR";
      opt.FilterDynamicsMatrices();
      std::vector<double> fr;
      for (int t = 0; t < T; t++) fr.push_back(opt.A[t](3, 2));
      print_vec("A_fir_32", fr);
      std::vector<double> raw;
      for (int t = 0; t < T; t++) raw.push_back(keep[t](3, 2));
      print_vec("A_raw_32", raw);
      std::vector<double> raw21;
      for (int t = 0; t < T; t++) raw21.push_back(keep[t](2, 1));
      print_vec("A_raw_21", raw21);
      opt.A = keep;
      opt.filteringMethod = "none";
    }
    opt.lambda = 0.1;
    auto Uopt = opt.Optimise(&start, U0, 3, 1, T);
    std::vector<double> uo;
    for (auto& u : Uopt) uo.push_back(u(0, 0));
    print_vec("U_opt", uo);
    std::printf("\"initial_cost\": %.17g, \"final_cost\": %.17g, \"iterations\": %d, \"cost_reduction\": %.17g}\n",
                opt.initial_cost, opt.new_cost, opt.num_iterations, opt.cost_reduction);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "shim_test: %s\n", e.what());
    return 1;
  }
  return 0;
}
