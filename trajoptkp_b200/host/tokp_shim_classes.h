// Second half of the host mirror (see tokp_shim.h): ModelTranslator index maps, Differentiator, KeypointGenerator,
// Optimiser / iLQR.  Every method body is marshalling + C-ABI calls; no arithmetic of the path runs on the host.
#pragma once
#include "tokp_shim.h"

// ---- ModelTranslator: state-vector <-> qpos/qvel maps and the task description ---------------------------------
// (InitModelTranslator ModelTranslator.cpp:18-96; ComputeStateDofAdrIndices :1548-1610; ReturnStateVector :882)
// one rigid body of the state vector (struct rigid_body, include/StdInclude.h:119-188): the active linear DoFs of its
// free joint become state elements after the robot joints (ComputeStateDofAdrIndices, ModelTranslator.cpp:1548-1610);
// angular elements (axis-angle read-out) are outside the supported subset
struct rigid_body {
  std::string name;
  bool active_linear_dof[3] = {false, false, false};
  bool active_angular_dof[3] = {false, false, false};
  double linear_jerk_threshold[3] = {0, 0, 0};
  double linear_vel_change_threshold[3] = {0, 0, 0};
};

class ModelTranslator {
 public:
  // Signature note: the reference's InitModelTranslator takes the task YAML path
  // (include/ModelTranslator/ModelTranslator.h:91) and reads these lists through yaml-cpp, which is not available here.
  // This overload adds what the manipulation tasks need on top of the joint-only form below: rigid bodies of the
  // state vector, residual targets that are vectors (residual_list[i].target[j]: `target` then holds them
  // concatenated while weight / weight_terminal have one entry per residual), and the bodies the residuals read.
  void InitModelTranslator(std::shared_ptr<MuJoCoHelper> helper, int task_kind, const std::vector<std::string>& joints,
                           const std::vector<std::string>& actuators, const std::vector<rigid_body>& bodies,
                           const std::vector<double>& target, const std::vector<double>& weight,
                           const std::vector<double>& weight_terminal, keypoint_method kp,
                           const std::vector<std::string>& aux_bodies) {
    std::vector<int> bj(helper->nbody + 8), jq(helper->nq + 64), jd(helper->nv + 64);
    tokp_model_get_int(helper->model, "body_jntadr", bj.data(), (int)bj.size());
    tokp_model_get_int(helper->model, "jnt_qposadr", jq.data(), (int)jq.size());
    tokp_model_get_int(helper->model, "jnt_dofadr", jd.data(), (int)jd.size());
    size_t nj = joints.size();
    // thresholds: the robot's come first in kp, the bodies' are appended in state order
    if (kp.jerk_thresholds.size() != nj) kp.jerk_thresholds.assign(nj, 0.0);
    if (kp.velocity_change_thresholds.size() != nj) kp.velocity_change_thresholds.assign(nj, 0.0);
    std::vector<int> bq, bd;
    for (const auto& b : bodies) {
      int id = helper->BodyId(b.name);
      if (id < 0) throw std::runtime_error("body not in model: " + b.name);
      if (b.active_angular_dof[0] || b.active_angular_dof[1] || b.active_angular_dof[2])
        throw std::runtime_error("angular rigid-body state elements are not supported");
      for (int k = 0; k < 3; k++)
        if (b.active_linear_dof[k]) {
          bq.push_back(jq[bj[id]] + k);
          bd.push_back(jd[bj[id]] + k);
          kp.jerk_thresholds.push_back(b.linear_jerk_threshold[k]);
          kp.velocity_change_thresholds.push_back(b.linear_vel_change_threshold[k]);
        }
    }
    keypoint_method robot_kp = kp;
    robot_kp.jerk_thresholds.resize(nj);
    robot_kp.velocity_change_thresholds.resize(nj);
    InitModelTranslator(helper, task_kind, joints, actuators, target, weight, weight_terminal, robot_kp);
    qpos_adr.insert(qpos_adr.end(), bq.begin(), bq.end());
    state_dof_adr_indices.insert(state_dof_adr_indices.end(), bd.begin(), bd.end());
    dof = (int)qpos_adr.size();
    current_kp = kp;
    aux.clear();
    aux_names = aux_bodies;
    for (const auto& n : aux_bodies) {
      int id = helper->BodyId(n);
      if (id < 0) throw std::runtime_error("body not in model: " + n);
      aux.push_back(id);
    }
  }
  // joint / actuator names in YAML order (robots[].jointNames / actuatorNames), residual terms flattened with
  // resid_dimension replication (FileHandler.cpp:276-287)
  void InitModelTranslator(std::shared_ptr<MuJoCoHelper> helper, int task_kind, const std::vector<std::string>& joints,
                           const std::vector<std::string>& actuators, const std::vector<double>& target,
                           const std::vector<double>& weight, const std::vector<double>& weight_terminal,
                           const keypoint_method& kp) {
    MuJoCo_helper = helper;
    kind = task_kind;
    std::vector<int> jq(helper->nq + 64), jd(helper->nv + 64);
    int nj = tokp_model_get_int(helper->model, "jnt_qposadr", jq.data(), (int)jq.size());
    tokp_model_get_int(helper->model, "jnt_dofadr", jd.data(), (int)jd.size());
    for (const auto& n : joints) {
      int id = helper->JointId(n);
      if (id < 0 || id >= nj) throw std::runtime_error("joint not in model: " + n);
      qpos_adr.push_back(jq[id]);
      state_dof_adr_indices.push_back(jd[id]);
    }
    for (const auto& n : actuators) {
      int id = helper->ActuatorId(n);
      if (id < 0) throw std::runtime_error("actuator not in model: " + n);
      ctrl_adr.push_back(id);
    }
    dof = (int)joints.size(); num_ctrl = (int)actuators.size();
    resid_target = target; resid_weight = weight; resid_weight_terminal = weight_terminal;
    current_kp = kp;
    if ((int)current_kp.jerk_thresholds.size() != dof) current_kp.jerk_thresholds.assign(dof, 0.0);
    if ((int)current_kp.velocity_change_thresholds.size() != dof) current_kp.velocity_change_thresholds.assign(dof, 0.0);
  }
  tokp_task_desc Desc() const {
    tokp_task_desc d;
    d.kind = kind; d.n = dof; d.m = num_ctrl;
    d.qpos_adr = qpos_adr.data(); d.dof_adr = state_dof_adr_indices.data(); d.ctrl_adr = ctrl_adr.data();
    d.num_resid = (int)resid_weight.size();
    d.resid_target = resid_target.data(); d.resid_weight = resid_weight.data();
    d.resid_weight_terminal = resid_weight_terminal.data();
    d.kp_method = tokp_kp_code(current_kp.name); d.min_N = current_kp.min_N; d.max_N = current_kp.max_N;
    d.jerk_thresholds = current_kp.jerk_thresholds.data();
    d.velchange_thresholds = current_kp.velocity_change_thresholds.data();
    d.iterative_error_threshold = current_kp.iterative_error_threshold;
    d.num_target = resid_target.size() != resid_weight.size() ? (int)resid_target.size() : 0;
    d.num_aux = (int)aux.size(); d.aux = aux.empty() ? nullptr : aux.data();
    return d;
  }
  // ReturnStateVector (ModelTranslator.cpp:882-897): [pos(n); vel(n)]
  MatrixXd ReturnStateVector(mjData* d) const {
    MatrixXd x(2 * dof, 1);
    for (int i = 0; i < dof; i++) {
      x(i, 0) = MuJoCo_helper->qpos(d)[qpos_adr[i]];
      x(i + dof, 0) = MuJoCo_helper->qvel(d)[state_dof_adr_indices[i]];
    }
    return x;
  }
  // StateIndexToQposIndex returns a DoF address despite its name (ModelTranslator.cpp:1707)
  int StateIndexToQposIndex(int state_index) const { return state_dof_adr_indices[state_index]; }
  MatrixXd ReturnControlVector(mjData* d) const {
    MatrixXd u(num_ctrl, 1);
    for (int i = 0; i < num_ctrl; i++) u(i, 0) = MuJoCo_helper->ctrl(d)[ctrl_adr[i]];
    return u;
  }
  // ReturnStateVectorQuaternions (ModelTranslator.cpp:899-950): the supported state vectors hold joint DoFs only, so
  // dof_quat == dof and the quaternion form is the plain state vector
  MatrixXd ReturnStateVectorQuaternions(mjData* d) const { return ReturnStateVector(d); }
  // ReturnPositionVector / ReturnVelocityVector (ModelTranslator.cpp:1014, :1173)
  MatrixXd ReturnPositionVector(mjData* d) const {
    MatrixXd p(dof, 1);
    for (int i = 0; i < dof; i++) p(i, 0) = MuJoCo_helper->qpos(d)[qpos_adr[i]];
    return p;
  }
  MatrixXd ReturnVelocityVector(mjData* d) const {
    MatrixXd v(dof, 1);
    for (int i = 0; i < dof; i++) v(i, 0) = MuJoCo_helper->qvel(d)[state_dof_adr_indices[i]];
    return v;
  }
  // SetControlVector (:989), SetPositionVector / SetVelocityVector (:1466), SetStateVector: scatter with a size check
  // (the reference prints and returns false on a size mismatch)
  bool SetControlVector(const MatrixXd& u, mjData* d) const {
    if (u.rows() != num_ctrl) { std::cerr << "ERROR: control vector size does not match the size of the control vector in the state vector list!" << std::endl; return false; }
    for (int i = 0; i < num_ctrl; i++) MuJoCo_helper->ctrl(d)[ctrl_adr[i]] = u(i, 0);
    return true;
  }
  bool SetPositionVector(const MatrixXd& p, mjData* d) const {
    if (p.rows() != dof) { std::cerr << "ERROR: position vector size does not match the size of the state vector list!" << std::endl; return false; }
    for (int i = 0; i < dof; i++) MuJoCo_helper->qpos(d)[qpos_adr[i]] = p(i, 0);
    return true;
  }
  bool SetVelocityVector(const MatrixXd& v, mjData* d) const {
    if (v.rows() != dof) { std::cerr << "ERROR: velocity vector size does not match the size of the state vector list!" << std::endl; return false; }
    for (int i = 0; i < dof; i++) MuJoCo_helper->qvel(d)[state_dof_adr_indices[i]] = v(i, 0);
    return true;
  }
  bool SetStateVector(const MatrixXd& x, mjData* d) const {
    if (x.rows() != 2 * dof) { std::cerr << "ERROR: state vector size does not match the size of the state vector list!" << std::endl; return false; }
    for (int i = 0; i < dof; i++) {
      MuJoCo_helper->qpos(d)[qpos_adr[i]] = x(i, 0);
      MuJoCo_helper->qvel(d)[state_dof_adr_indices[i]] = x(i + dof, 0);
    }
    return true;
  }
  // ReturnControlLimits (:971-987): [lo_0, hi_0, lo_1, hi_1, ...] read from actuator_ctrlrange (MuJoCoHelper.cpp:262-264)
  MatrixXd ReturnControlLimits() const {
    std::vector<double> r(2 * (size_t)MuJoCo_helper->nu + 2);
    tokp_model_get_real(MuJoCo_helper->model, "actuator_ctrlrange", r.data(), (int)r.size());
    MatrixXd lim(2 * num_ctrl, 1);
    for (int i = 0; i < num_ctrl; i++) { lim(2 * i, 0) = r[2 * ctrl_adr[i]]; lim(2 * i + 1, 0) = r[2 * ctrl_adr[i] + 1]; }
    return lim;
  }
  // MT::Residuals overrides (Acrobot.cpp:26-55, Walker.cpp:59-90) and MT::CostFunction (ModelTranslator.cpp:314-327) on
  // the host, for harness code that scores recorded states (SingleMPCRun's final-cost replay, GenTestingData.cpp:565-585);
  // the optimiser itself evaluates them on the device
  void Residuals(mjData* d, MatrixXd& residuals) const {
    const double* q = MuJoCo_helper->qpos(d);
    const double* v = MuJoCo_helper->qvel(d);
    const double* u = MuJoCo_helper->ctrl(d);
    const std::vector<double>& t = resid_target;
    if ((int)residuals.rows() != (int)resid_weight.size()) residuals.resize((int)resid_weight.size(), 1);
    int k = 0;
    if (kind == TOKP_TASK_BOXSWEEP) {
      // BoxSweep::Residuals (BoxSweep.cpp:101-139): goal body to its target (x, y), its velocity (x, y), gripper to the
      // goal body minus the reach distance; the goal body's linear elements follow the robot joints in the state
      if (aux_names.size() < 2) throw std::runtime_error("box_sweep needs aux bodies {goal, franka_gripper}");
      int nj = dof - 3;
      double gp[3] = {q[qpos_adr[nj]], q[qpos_adr[nj + 1]], q[qpos_adr[nj + 2]]};
      double gv[2] = {v[state_dof_adr_indices[nj]], v[state_dof_adr_indices[nj + 1]]};
      residuals(k++, 0) = std::sqrt(std::pow(gp[0] - t[0], 2) + std::pow(gp[1] - t[1], 2));
      residuals(k++, 0) = std::sqrt(std::pow(gv[0] - t[2], 2) + std::pow(gv[1] - t[3], 2));
      double ee[3];
      MuJoCo_helper->BodyXpos(aux_names[1], d, ee);
      double dist = std::sqrt(std::pow(ee[0] - gp[0], 2) + std::pow(ee[1] - gp[1], 2) + std::pow(ee[2] - gp[2], 2));
      residuals(k++, 0) = dist - t[4];
      return;
    }
    if (kind == TOKP_TASK_ACROBOT) {
      residuals(k++, 0) = q[qpos_adr[0]] - t[0];
      residuals(k++, 0) = q[qpos_adr[1]] - t[1];
      residuals(k++, 0) = v[state_dof_adr_indices[0]] - t[2];
      residuals(k++, 0) = v[state_dof_adr_indices[1]] - -t[3];  // the double minus is the reference's (Acrobot.cpp:45)
      residuals(k++, 0) = u[ctrl_adr[0]] - t[4];
    } else if (kind == TOKP_TASK_WALKER) {
      residuals(k++, 0) = q[qpos_adr[0]] - t[0];                // body height
      residuals(k++, 0) = q[qpos_adr[2]] - t[1];                // body rotation
      residuals(k++, 0) = v[state_dof_adr_indices[1]] - t[2];   // forward velocity
      for (int i = 0; i < num_ctrl; i++, k++) residuals(k, 0) = u[ctrl_adr[i]] - t[3 + i];
    } else {
      throw std::runtime_error("residuals not implemented for this model");
    }
    if (k != (int)resid_weight.size()) throw std::runtime_error("Error: Residuals size mismatch");
  }
  double CostFunction(const MatrixXd& residuals, bool terminal) const {
    double cost = 0.0;
    for (size_t i = 0; i < resid_weight.size(); i++)
      cost += (terminal ? resid_weight_terminal[i] : resid_weight[i]) * std::pow(residuals((int)i, 0), 2);
    return cost;
  }

  std::shared_ptr<MuJoCoHelper> MuJoCo_helper;
  int kind = 0, dof = 0, num_ctrl = 0;
  std::vector<int> qpos_adr, state_dof_adr_indices, ctrl_adr, aux;
  std::vector<std::string> aux_names;
  std::vector<double> resid_target, resid_weight, resid_weight_terminal;
  keypoint_method current_kp;
};

// Device arena shared by Differentiator / KeypointGenerator / Optimiser of one problem (replaces the per-thread
// fd_data[] scratch and the T+1 saved mjData allocations, MuJoCoHelper.cpp:696-703,918-921)
struct TrajHandle {
  TrajHandle(std::shared_ptr<ModelTranslator> mt, int horizon, int batch = 1) : T(horizon), B(batch), mt_(mt) {
    tokp_task_desc d = mt->Desc();
    tokp_shim::check(tokp_traj_create(mt->MuJoCo_helper->model, &d, horizon, batch, &h), "tokp_traj_create");
  }
  ~TrajHandle() { if (h) tokp_traj_destroy(h); }
  TrajHandle(const TrajHandle&) = delete;
  tokp_traj* h = nullptr;
  int T, B;
  std::shared_ptr<ModelTranslator> mt_;
};

// ---- Differentiator (include/Differentiator.h:15-37) -------------------------------------------------------------
class Differentiator {
 public:
  Differentiator(std::shared_ptr<ModelTranslator> model_translator, std::shared_ptr<MuJoCoHelper> MuJoCo_helper,
                 std::shared_ptr<TrajHandle> traj)
      : model_translator(model_translator), MuJoCo_helper(MuJoCo_helper), traj(traj) {}
  // Differentiator::DynamicsDerivatives (Differentiator.cpp:8-462): writes columns {c, c+n} of A and {c | c<m} of B
  // for c in cols at saved_systems_state_list[data_index]; thread_id is accepted and ignored (no per-thread scratch).
  void DynamicsDerivatives(MatrixXd& A, MatrixXd& B, const std::vector<int>& cols, int data_index, int /*thread_id*/,
                           bool central_diff, double eps) {
    // the saved state of this index is pushed first, exactly what the reference reads (Differentiator.cpp:62-66)
    const auto& s = MuJoCo_helper->saved_systems_state_list.at(data_index)->v;
    tokp_shim::check(tokp_traj_set_snapshot(traj->h, 0, data_index, s.data()), "set_snapshot");
    tokp_shim::check(tokp_dynamics_derivs_cols(traj->h, 0, data_index, cols.data(), (int)cols.size(), central_diff ? 1 : 0,
                                               eps, A.data(), A.rows(), B.data(), B.rows()), "DynamicsDerivatives");
  }
  // Differentiator::ResidualDerivatives (Differentiator.cpp:464-663): r_x[j] is nx x 1 and r_u[j] is m x 1 for every
  // residual j, finite differences of MT::Residuals at saved_systems_state_list[data_index]
  void ResidualDerivatives(std::vector<MatrixXd>& r_x, std::vector<MatrixXd>& r_u, int data_index, int /*tid*/,
                           bool central_diff, double eps) {
    const int nx = 2 * model_translator->dof, m = model_translator->num_ctrl;
    const int nr = (int)model_translator->resid_weight.size();
    const auto& s = MuJoCo_helper->saved_systems_state_list.at(data_index)->v;
    tokp_shim::check(tokp_traj_set_snapshot(traj->h, 0, data_index, s.data()), "set_snapshot");
    std::vector<double> rx((size_t)nx * nr), ru((size_t)(m ? m : 1) * nr);
    tokp_shim::check(tokp_residual_derivs(traj->h, 0, data_index, central_diff ? 1 : 0, eps, rx.data(), ru.data()),
                     "ResidualDerivatives");
    r_x.assign(nr, MatrixXd(nx, 1));
    r_u.assign(nr, MatrixXd(m, 1));
    for (int j = 0; j < nr; j++) {
      for (int c = 0; c < nx; c++) r_x[j](c, 0) = rx[(size_t)c * nr + j];
      for (int c = 0; c < m; c++) r_u[j](c, 0) = ru[(size_t)c * nr + j];
    }
  }
  std::shared_ptr<ModelTranslator> model_translator;
  std::shared_ptr<MuJoCoHelper> MuJoCo_helper;
  std::shared_ptr<TrajHandle> traj;
};

// ---- KeypointGenerator (include/KeyPointGenerator.h:56-110) ------------------------------------------------------
class KeypointGenerator {
 public:
  KeypointGenerator(std::shared_ptr<Differentiator> differentiator, std::shared_ptr<MuJoCoHelper> /*helper*/, int _dof,
                    int _horizon)
      : dof(_dof), horizon(_horizon), differentiator_(differentiator) {
    last_percentages.assign(dof, 0.0);
  }
  void SetKeypointMethod(keypoint_method m) {
    method_ = m;
    tokp_shim::check(tokp_traj_set_keypoint_method(differentiator_->traj->h, tokp_kp_code(m.name), m.min_N, m.max_N,
                                                   m.jerk_thresholds.empty() ? nullptr : m.jerk_thresholds.data(),
                                                   m.velocity_change_thresholds.empty() ? nullptr : m.velocity_change_thresholds.data(),
                                                   m.iterative_error_threshold), "SetKeypointMethod");
  }
  keypoint_method ReturnCurrentKeypointMethod() { return method_; }
  void ResetCache() {}  // computed_keypoints cache lives on the device and is rebuilt per call
  // GenerateKeyPoints (KeyPointGenerator.cpp:76-135): trajectory_states are X_old[0..T]; A, B are only touched by
  // iterative_error (filled at the key-points it evaluates)
  void GenerateKeyPoints(const std::vector<MatrixXd>& trajectory_states, std::vector<MatrixXd>& A, std::vector<MatrixXd>& B) {
    tokp_traj* h = differentiator_->traj->h;
    int nx = 2 * dof;
    std::vector<double> X((size_t)(horizon + 1) * nx, 0.0);
    for (size_t t = 0; t < trajectory_states.size() && t <= (size_t)horizon; t++)
      std::memcpy(&X[t * nx], trajectory_states[t].data(), sizeof(double) * nx);
    tokp_shim::check(tokp_traj_set(h, 0, "X_old", X.data(), (int)X.size()), "set X_old");
    std::vector<int> counts(horizon), idx((size_t)horizon * 2 * dof);
    tokp_shim::check(tokp_keypoints(h, counts.data(), idx.data(), last_percentages.data()), "GenerateKeyPoints");
    keypoints.assign(horizon, {});
    for (int t = 0; t < horizon; t++) keypoints[t].assign(&idx[(size_t)t * 2 * dof], &idx[(size_t)t * 2 * dof] + counts[t]);
    if (method_.name == "iterative_error") PullAB(A, B);
  }
  // InterpolateDerivatives (KeyPointGenerator.cpp:840-954) on caller-owned A/B and an arbitrary key-point table
  void InterpolateDerivatives(const std::vector<std::vector<int>>& keyPoints, int T, std::vector<MatrixXd>& A,
                              std::vector<MatrixXd>& B, std::vector<std::vector<MatrixXd>>& /*r_x*/,
                              std::vector<std::vector<MatrixXd>>& /*r_u*/, bool /*residual_derivs*/, int num_ctrl) {
    tokp_traj* h = differentiator_->traj->h;
    int nx = 2 * dof;
    std::vector<int> counts(T), idx;
    for (int t = 0; t < T; t++) { counts[t] = (int)keyPoints[t].size(); idx.insert(idx.end(), keyPoints[t].begin(), keyPoints[t].end()); }
    idx.push_back(0);
    tokp_shim::check(tokp_traj_set_keypoints(h, 0, counts.data(), idx.data()), "set_keypoints");
    std::vector<double> a((size_t)T * nx * nx), b((size_t)T * nx * num_ctrl);
    for (int t = 0; t < T; t++) {
      std::memcpy(&a[(size_t)t * nx * nx], A[t].data(), sizeof(double) * nx * nx);
      if (num_ctrl) std::memcpy(&b[(size_t)t * nx * num_ctrl], B[t].data(), sizeof(double) * nx * num_ctrl);
    }
    tokp_shim::check(tokp_traj_set(h, 0, "A", a.data(), (int)a.size()), "set A");
    if (num_ctrl) tokp_shim::check(tokp_traj_set(h, 0, "B", b.data(), (int)b.size()), "set B");
    tokp_shim::check(tokp_interpolate(h), "InterpolateDerivatives");
    PullAB(A, B);
  }
  void PullAB(std::vector<MatrixXd>& A, std::vector<MatrixXd>& B) {
    tokp_traj* h = differentiator_->traj->h;
    int nx = 2 * dof, T = horizon, m = B.empty() ? 0 : B[0].cols();
    std::vector<double> a((size_t)T * nx * nx), b((size_t)T * nx * (m ? m : 1));
    tokp_traj_get(h, 0, "A", a.data(), (int)a.size());
    if (m) tokp_traj_get(h, 0, "B", b.data(), (int)b.size());
    for (int t = 0; t < T && t < (int)A.size(); t++) {
      std::memcpy(A[t].data(), &a[(size_t)t * nx * nx], sizeof(double) * nx * nx);
      if (m) std::memcpy(B[t].data(), &b[(size_t)t * nx * m], sizeof(double) * nx * m);
    }
  }
  int dof, horizon;
  std::vector<std::vector<int>> keypoints;
  std::vector<double> last_percentages;

 private:
  std::shared_ptr<Differentiator> differentiator_;
  keypoint_method method_;
};
