// C++ host mirror of the reference's class interfaces for the hot path, implemented ONLY on top of the C ABI in
// include/tokp.h (libtokp.so).  A TrajOptKP maintainer replaces the bodies of the cited methods with these; names,
// argument meaning and ownership follow the reference:
//   Differentiator            include/Differentiator.h:15-37          src/Differentiator/Differentiator.cpp:8-462
//   KeypointGenerator         include/KeyPointGenerator.h:34-110      src/KeyPointGenerator/KeyPointGenerator.cpp:76-135,840-954
//   Optimiser / iLQR          include/Optimiser/Optimiser.h:49-242    src/Optimiser/iLQR.cpp:202-410
//   ModelTranslator (maps)    include/ModelTranslator/ModelTranslator.h:91-177
//   MuJoCoHelper (buffers)    include/MuJoCoHelper.h:22-31,118-123    src/PhysicsSimulators/MuJoCoHelper.cpp:696-741,877-937
// Differences forced by the boundary: `mjData*` becomes `SystemState*` (the 8-field snapshot of CpMjData,
// MuJoCoHelper.cpp:732-741); Eigen's MatrixXd becomes the minimal column-major stand-in below when Eigen is absent
// (define TOKP_SHIM_USE_EIGEN to use the real one); errors throw std::runtime_error instead of exit(1).
#pragma once
#include <cmath>
#include <cstring>
#include <iostream>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "tokp.h"

#ifdef TOKP_SHIM_USE_EIGEN
#include <Eigen/Dense>
using Eigen::MatrixXd;
#else
// column-major dynamic matrix with the handful of members the path uses (Eigen::MatrixXd's default layout)
class MatrixXd {
 public:
  MatrixXd() = default;
  MatrixXd(int r, int c) : r_(r), c_(c), d_((size_t)r * c, 0.0) {}
  int rows() const { return r_; }
  int cols() const { return c_; }
  double& operator()(int i, int j) { return d_[(size_t)j * r_ + i]; }
  double operator()(int i, int j) const { return d_[(size_t)j * r_ + i]; }
  double* data() { return d_.data(); }
  const double* data() const { return d_.data(); }
  void setZero() { std::fill(d_.begin(), d_.end(), 0.0); }
  void resize(int r, int c) { r_ = r; c_ = c; d_.assign((size_t)r * c, 0.0); }

 private:
  int r_ = 0, c_ = 0;
  std::vector<double> d_;
};
#endif

namespace tokp_shim {

inline void check(int rc, const char* what) {
  if (rc != TOKP_OK) throw std::runtime_error(std::string(what) + ": " + tokp_last_error());
}

// mjData stand-in: [time | qpos | qvel | qacc | qacc_warmstart | qfrc_applied | xfrc_applied | ctrl]
struct SystemState {
  std::vector<double> v;
};

}  // namespace tokp_shim

// struct keypoint_method, include/KeyPointGenerator.h:34-43
struct keypoint_method {
  std::string name = "set_interval";
  bool auto_adjust = false;
  int min_N = 1;
  int max_N = 10;
  std::vector<double> jerk_thresholds;
  std::vector<double> accell_thresholds;
  double iterative_error_threshold = 0.0;
  std::vector<double> velocity_change_thresholds;
};

inline int tokp_kp_code(const std::string& name) {
  if (name == "set_interval" || name == "setInterval") return TOKP_KP_SET_INTERVAL;
  if (name == "adaptive_jerk") return TOKP_KP_ADAPTIVE_JERK;
  if (name == "adaptive_accel") return TOKP_KP_ADAPTIVE_ACCEL;
  if (name == "iterative_error") return TOKP_KP_ITERATIVE_ERROR;
  if (name == "velocity_change") return TOKP_KP_VELOCITY_CHANGE;
  throw std::runtime_error("unknown keypoint method " + name);  // reference: cerr + exit(1), KeyPointGenerator.cpp:130-133
}

// ---- MuJoCoHelper: owns the model and the saved_systems_state_list (include/MuJoCoHelper.h:118-123) --------------
class MuJoCoHelper {
 public:
  using mjData = tokp_shim::SystemState;
  // InitSimulator(timestep, file_path, use_plugins), MuJoCoHelper.cpp:877-923
  void InitSimulator(double timestep, const char* file_path, bool /*use_plugins*/ = false) {
    tokp_shim::check(tokp_model_compile(file_path, timestep, &model), "InitSimulator");
    int d[11];
    tokp_model_dims(model, d);
    nq = d[0]; nv = d[1]; nu = d[2]; nbody = d[3];
    S = tokp_snapshot_size(model);
    master_reset_data.reset(new mjData{std::vector<double>(S, 0.0)});
    std::vector<double> q0(nq);
    tokp_model_get_real(model, "qpos0", q0.data(), nq);
    std::memcpy(master_reset_data->v.data() + 1, q0.data(), sizeof(double) * nq);
    main_data.reset(new mjData(*master_reset_data));
  }
  ~MuJoCoHelper() { if (model) tokp_model_destroy(model); }
  // CopySystemState / CpMjData, MuJoCoHelper.cpp:705-741
  bool CopySystemState(mjData* dest, const mjData* src) { dest->v = src->v; return true; }
  bool AppendSystemStateToEnd(const mjData* d) { saved_systems_state_list.emplace_back(new mjData(*d)); return true; }
  bool CheckIfDataIndexExists(int i) const { return i >= 0 && i < (int)saved_systems_state_list.size(); }
  bool ClearSystemStateList() { saved_systems_state_list.clear(); return true; }
  double ReturnModelTimeStep() const { double o[8]; tokp_model_get_real(model, "opt_real", o, 8); return o[0]; }
  // name -> index (mj_name2id call sites MuJoCoHelper.cpp:36-438)
  int JointId(const std::string& n) const { return tokp_model_name2id(model, TOKP_OBJ_JOINT, n.c_str()); }
  int ActuatorId(const std::string& n) const { return tokp_model_name2id(model, TOKP_OBJ_ACTUATOR, n.c_str()); }
  int BodyId(const std::string& n) const { return tokp_model_name2id(model, TOKP_OBJ_BODY, n.c_str()); }
  // GetBodyPoseQuatViaXpos (MuJoCoHelper.cpp:400-410) after mj_kinematics: world position of a body at d's qpos
  void BodyXpos(const std::string& n, mjData* d, double pos[3]) const {
    double quat[4];
    tokp_shim::check(tokp_model_body_pose(model, qpos(d), BodyId(n), pos, quat), "BodyXpos");
  }
  double* qpos(mjData* d) const { return d->v.data() + 1; }
  double* qvel(mjData* d) const { return d->v.data() + 1 + nq; }
  double* ctrl(mjData* d) const { return d->v.data() + S - nu; }

  tokp_model* model = nullptr;
  int nq = 0, nv = 0, nu = 0, nbody = 0, S = 0;
  std::vector<std::unique_ptr<mjData>> saved_systems_state_list;
  std::unique_ptr<mjData> main_data, master_reset_data;
};
using mjData = MuJoCoHelper::mjData;
