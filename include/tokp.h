/* tokp - C ABI of the B200-native key-point FD-derivative + iLQR path.
 *
 * The reference (DMackRus/TrajOptKP) has no FFI layer: the path sits behind C++ classes
 * (include/Differentiator.h:17-24, include/KeyPointGenerator.h:56-110, include/Optimiser/Optimiser.h:49-242,
 *  include/Optimiser/iLQR.h, include/MuJoCoHelper.h:118-123).  Each entry point below names the reference
 * method whose body it replaces; the C++ shim classes in trajoptkp_b200/host/ keep those class
 * signatures and call only these functions.  All pointers are HOST pointers owned by the caller;
 * all matrices are column-major with leading dimension = rows (Eigen's default layout).
 * Every function returns TOKP_OK (0) or an error code; tokp_last_error() gives the message.
 * There is no CPU fallback: without a CUDA device every compute entry point returns TOKP_ERR_CUDA.
 */
#ifndef TOKP_H_
#define TOKP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tokp_model tokp_model; /* compiled mjModel equivalent (host + device copy)            */
typedef struct tokp_traj tokp_traj;   /* batch of B trajectories: device arena + iLQR state          */

enum {
  TOKP_OK = 0,
  TOKP_ERR_ARG = 1,         /* bad argument / size                                                    */
  TOKP_ERR_MODEL = 2,       /* MJCF outside the supported subset, parse error                         */
  TOKP_ERR_CUDA = 3,        /* CUDA runtime error or no device                                        */
  TOKP_ERR_UNSUPPORTED = 4, /* feature outside the hot-path scope (e.g. non-zero applied forces)      */
  TOKP_ERR_STATE = 5        /* call sequence error (e.g. derivatives before a rollout)                */
};

/* object types for tokp_model_name2id (mj_name2id equivalents used by MuJoCoHelper.cpp:36-438) */
enum { TOKP_OBJ_BODY = 1, TOKP_OBJ_JOINT = 3, TOKP_OBJ_GEOM = 5, TOKP_OBJ_SITE = 6, TOKP_OBJ_ACTUATOR = 19 };

/* task kinds: which MT::Residuals override runs on the device
 * (src/ModelTranslator/Acrobot.cpp:26-55, Walker.cpp:59-90) */
enum { TOKP_TASK_ACROBOT = 0, TOKP_TASK_WALKER = 1, TOKP_TASK_REACHING = 2 /* pandaReaching, src/ModelTranslator/Reaching.cpp:28-60 */,
       TOKP_TASK_BOXSWEEP = 3 /* BoxSweep, src/ModelTranslator/BoxSweep.cpp:101-139 */,
       TOKP_TASK_PUSH = 4 /* TwoDPushing (no / light / heavy clutter), src/ModelTranslator/TwoDPushing.cpp:291-356;
                             aux = {goal body, franka_gripper body}; targets = goal_pos (x, y), goal_vel (x, y), then
                             per obstacle its start (x, y), then joint_5_velocity (1), reach (1) */ };

/* key-point methods (include/KeyPointGenerator.h:34-43) */
enum {
  TOKP_KP_SET_INTERVAL = 0,
  TOKP_KP_ADAPTIVE_JERK = 1,
  TOKP_KP_ADAPTIVE_ACCEL = 2,
  TOKP_KP_ITERATIVE_ERROR = 3,
  TOKP_KP_VELOCITY_CHANGE = 4
};

/* Optimiser::filteringMethod "none" | "low_pass" | "FIR" (include/Optimiser/Optimiser.h:173) */
enum { TOKP_FILTER_NONE = 0, TOKP_FILTER_LOW_PASS = 1, TOKP_FILTER_FIR = 2 };

/* mjSTAGE_* as used by mj_stepSkip (Differentiator.cpp:118,253,357) */
enum { TOKP_STAGE_NONE = 0, TOKP_STAGE_POS = 1, TOKP_STAGE_VEL = 2 };

/* What ModelTranslator::InitModelTranslator derives from the task YAML
 * (src/ModelTranslator/ModelTranslator.cpp:18-96, stateVectorList include/StdInclude.h:110-188):
 * the optimiser state is [pos(n); vel(n)], controls u(m). */
typedef struct tokp_task_desc {
  int kind;                             /* TOKP_TASK_*                                                */
  int n;                                /* optimiser DoFs (current_state_vector.dof)                  */
  int m;                                /* controls (num_ctrl)                                        */
  const int* qpos_adr;                  /* [n] qpos address of each state position                    */
  const int* dof_adr;                   /* [n] dof address (ModelTranslator::StateIndexToQposIndex)   */
  const int* ctrl_adr;                  /* [m] actuator index of each control                         */
  int num_resid;                        /* residual_list.size()                                       */
  const double* resid_target;           /* [num_resid] residual_list[i].target[0]                     */
  const double* resid_weight;           /* [num_resid]                                                */
  const double* resid_weight_terminal;  /* [num_resid]                                                */
  int kp_method;                        /* TOKP_KP_*                                                  */
  int min_N, max_N;
  const double* jerk_thresholds;        /* [n]                                                        */
  const double* velchange_thresholds;   /* [n]                                                        */
  double iterative_error_threshold;
  /* Residuals whose targets are vectors (BoxSweep: goal_pos (x, y), goal_vel (x, y), reach (d)) and residuals that
   * read body poses.  num_target = 0 means num_resid; resid_target then holds num_target scalars, the residuals'
   * target vectors concatenated (residual_list[i].target[j], include/StdInclude.h:82-88).  aux[num_aux] = model
   * ids the residual needs: BoxSweep {goal body id, franka_gripper body id}. */
  int num_target;
  int num_aux;
  const int* aux;
} tokp_task_desc;

const char* tokp_last_error(void);
const char* tokp_version(void);

/* ---- model (MuJoCoHelper::InitSimulator, src/PhysicsSimulators/MuJoCoHelper.cpp:877-923) ---------- */
/* TOKP_ERR_MODEL for anything the step does not model and mj_loadXML would have accepted or rejected itself: collision
 * pairs without a narrow-phase function, condim other than 1 / 3, joint frictionloss / stiffness, <tendon>, <equality>,
 * OBJ face indices outside the vertex list, integrators / cones / solvers other than Euler / pyramidal / Newton.
 * The handle is uploaded to the CUDA device that is current at its first device use and is valid on that device
 * only: using it with another current device returns TOKP_ERR_STATE (compile the model once per device). */
int tokp_model_compile(const char* mjcf_path, double timestep, tokp_model** out);
int tokp_model_compile_string(const char* mjcf_xml, const char* base_dir, double timestep, tokp_model** out);
void tokp_model_destroy(tokp_model* m);
/* dims = [nq, nv, nu, nbody, njnt, ngeom, npair, nmeshvert, maxcontact, maxefc, treedepth] */
int tokp_model_dims(const tokp_model* m, int dims[11]);
int tokp_model_name2id(const tokp_model* m, int objtype, const char* name); /* -1 when absent        */
/* serialised model blob; returns the size in bytes (call with buf = NULL to query) or -1 */
int64_t tokp_model_blob(const tokp_model* m, void* buf, int64_t cap);
/* copy a named real / int field (names: trajoptkp_b200/csrc/model.h); returns count, -count if cap too small, -1e6 unknown */
int tokp_model_get_real(const tokp_model* m, const char* field, double* out, int cap);
int tokp_model_get_int(const tokp_model* m, const char* field, int32_t* out, int cap);
/* doubles per snapshot: [time | qpos nq | qvel nv | qacc nv | qacc_warmstart nv | qfrc_applied nv |
 * xfrc_applied 6*nbody | ctrl nu] - the 8 fields of MuJoCoHelper::CpMjData (MuJoCoHelper.cpp:732-741) */
int tokp_snapshot_size(const tokp_model* m);
/* mj_kinematics restricted to the chain world -> body (host): world position and orientation of a body at qpos[nq],
 * what MuJoCoHelper::GetBodyPoseQuatViaXpos (MuJoCoHelper.cpp:400-410) reads after mj_kinematics */
int tokp_model_body_pose(const tokp_model* m, const double* qpos, int body, double pos[3], double quat[4]);

/* ---- trajectory batch (Optimiser storage include/Optimiser/Optimiser.h:194-211; iLQR::Resize iLQR.cpp:37-200;
 *      saved_systems_state_list / fd_data include/MuJoCoHelper.h:118-123) --------------------------------- */
int tokp_traj_create(const tokp_model* m, const tokp_task_desc* task, int horizon, int batch, tokp_traj** out);
void tokp_traj_destroy(tokp_traj* tr);
/* per-trajectory residual targets (LoadTaskFromFile, src/FileHandler/FileHandler.cpp:471-578) */
int tokp_traj_set_targets(tokp_traj* tr, int b, const double* resid_target);
int tokp_traj_set_keypoint_method(tokp_traj* tr, int kp_method, int min_N, int max_N, const double* jerk_thr,
                                  const double* vel_thr, double iter_thr);
/* saved_systems_state_list[t] of trajectory b */
int tokp_traj_set_snapshot(tokp_traj* tr, int b, int t, const double* snapshot);
int tokp_traj_get_snapshot(tokp_traj* tr, int b, int t, double* snapshot);

/* batched plain stepping of `count` independent worlds: mj_step / mj_stepSkip(NONE) applied `nsteps` times.
 * solver_iterations < 0 selects the model's options (100, 1e-8). status[count] receives per-world warning bits. */
int tokp_step_worlds(const tokp_model* m, const double* snapshots_in, double* snapshots_out, int count, int nsteps,
                     int solver_iterations, double solver_tolerance, int* status);

/* the same followed by one mj_forward (ModelTranslator::CreateInitSetupControls, ModelTranslator.cpp:875-880, between the
 * plant step and Optimise in GenTestingData.cpp:189-205): the state is the one after `nsteps` steps, qacc and
 * qacc_warmstart are the constraint solver's solution at that state */
int tokp_forward_worlds(const tokp_model* m, const double* snapshots_in, double* snapshots_out, int count, int nsteps,
                        int solver_iterations, double solver_tolerance, int* status);

/* tokp_step_worlds without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  The kernel's scratch (input snapshots | output snapshots | status words) is a static device array of 4 MiB
 * (TOKP_STATIC_SCRATCH_BYTES) that every GPU loading the library holds until exit; a call needs
 * count * (2 * 8 * tokp_snapshot_size(m) + 4) bytes of it and returns TOKP_ERR_ARG beyond that (walker: 2614 worlds).
 * The model must already be on the device (any earlier tokp_traj_create / tokp_step_worlds on it; TOKP_ERR_STATE
 * otherwise).  Calls no cudaMalloc / cudaFree.  All calls share the array: calls that overlap on the GPU race. */
int tokp_step_worlds_static(const tokp_model* m, const double* snapshots_in, double* snapshots_out, int count, int nsteps,
                            int solver_iterations, double solver_tolerance, int* status);

/* iLQR::RolloutTrajectory (iLQR.cpp:202-254) for all B trajectories: snapshot0[B][S], U[B][T][m] -> cost[B] */
int tokp_rollout(tokp_traj* tr, const double* snapshot0, const double* U, double* cost);
/* pandaReaching::CreateInitOptimisationControls (src/ModelTranslator/Reaching.cpp:269-296): gravity-compensation
 * rollout from snapshot0[B][S]; control t = qfrc_bias of the state joints at step t (mj_forward), then mj_step.
 * U_out[B][T][m]; status[B] (optional) receives the per-world warning bits. */
int tokp_init_controls_gravcomp(tokp_traj* tr, const double* snapshot0, double* U_out, int* status);
/* PushBaseClass::JacobianEEControl (src/ModelTranslator/PushBaseClass.cpp:139-248) with its way-point builders
 * (EEWayPointsSetup :8-44, EEWayPointsPush :46-115, CreateAllEETransitPoints :117-137), driven in the order of
 * GenTestingData::GenDataOpenloopOptimisation (src/GenTestingData.cpp:177-220) for every trajectory of the batch:
 * `pre_steps` x mj_step, mj_forward, CreateInitSetupControls(setup_horizon) (TwoDPushing.cpp:225-258; 0 = BoxSweep.cpp:63-70,
 * no setup motion), snapshot0_out[B][S] <- state (what Optimise starts from), CreateInitOptimisationControls(T)
 * (BoxSweep.cpp:72-99, TwoDPushing.cpp:260-289) -> U_out[B][T][m].  snapshot_start[B][S] = the state
 * ModelTranslator::InitialiseSystemToStartState (ModelTranslator.cpp:1711-1781) leaves; goal_xy[B][2] =
 * residual_list[0].target; task aux = {pushed body, end-effector body}.  Optional: U_setup_out[B][setup_horizon][m],
 * status[B] (step warning bits; bit 6 = rank-deficient end-effector Jacobian). */
int tokp_init_controls_ee(tokp_traj* tr, const double* snapshot_start, const double* goal_xy, int pre_steps,
                          int setup_horizon, double* snapshot0_out, double* U_out, double* U_setup_out, int* status);
/* KeypointGenerator::GenerateKeyPoints (KeyPointGenerator.cpp:76-135) on the nominal X_old; optional outputs:
 * counts[B][T], indices[B][T][2n] (row lists, in reference order, duplicates kept), pct[B][n] */
int tokp_keypoints(tokp_traj* tr, int* counts, int* indices, double* pct);
/* override the device key-point rows of trajectory b (tests / Optimiser::SmoothDerivativesAtContact) */
int tokp_traj_set_keypoints(tokp_traj* tr, int b, const int* counts, const int* indices);
/* Optimiser::ComputeDynamicsDerivativesAtKeypoints (Optimiser.cpp:239-323): ONE batched launch over all
 * (b, t, column, +-eps) worlds of the current key-points; FD solver options (5 iterations, tol 0) */
int tokp_dynamics_derivs(tokp_traj* tr, int central_diff, double eps);
/* Differentiator::DynamicsDerivatives (Differentiator.cpp:8-462) exact drop-in for one (b, t, cols):
 * writes columns {c, c+n} of A (lda) and {c | c<m} of B (ldb), leaves the others untouched */
int tokp_dynamics_derivs_cols(tokp_traj* tr, int b, int t, const int* cols, int ncols, int central_diff, double eps,
                              double* A, int lda, double* B, int ldb);
/* KeypointGenerator::InterpolateDerivatives (KeyPointGenerator.cpp:840-954) */
int tokp_interpolate(tokp_traj* tr);
/* Optimiser::ComputeCostDerivatives (Optimiser.cpp:197-236): residual FD (Differentiator.cpp:464-663) for
 * t = 0..T and Gauss-Newton l_x, l_xx, l_u, l_uu (ModelTranslator.cpp:552-583), terminal at T-1 */
int tokp_cost_derivs(tokp_traj* tr, int central_diff, double eps);
/* Differentiator::ResidualDerivatives (Differentiator.cpp:464-663) exact drop-in for one (b, t), t = 0..T:
 * r_x[nx][nr] and r_u[m][nr] (entry [column][residual], i.e. the reference's r_x[j](column, 0)); r_u may be NULL */
int tokp_residual_derivs(tokp_traj* tr, int b, int t, int central_diff, double eps, double* r_x, double* r_u);
/* Optimiser::filteringMethod / lowPassACoefficient / setFIRFilter (include/Optimiser/Optimiser.h:151,173,219-220):
 * method TOKP_FILTER_*; fir == NULL keeps the current taps (default {0.1, 0.15, 0.5, 0.15, 0.1}); at most 16 taps */
int tokp_traj_set_filter(tokp_traj* tr, int method, double low_pass_a, const double* fir, int nfir);
/* Optimiser::FilterDynamicsMatrices (Optimiser.cpp:340-406): filters the velocity rows of A[0..T-1] along t in place;
 * TOKP_ERR_STATE when no method is set (the reference prints "Filtering method not recognised") */
int tokp_filter_dynamics(tokp_traj* tr);
/* Optimiser::GenerateDerivatives (Optimiser.cpp:80-108): key-points, dynamics derivatives, interpolation, cost
 * derivatives and (when a filter is set) FilterDynamicsMatrices, in the reference's order */
int tokp_generate_derivatives(tokp_traj* tr);
/* iLQR::BackwardsPassQuuRegularisation (iLQR.cpp:535-634): lambda[B] in; ok[B], delta_J[B] out */
int tokp_backward_pass(tokp_traj* tr, const double* lambda, int* ok, double* delta_J);
/* iLQR::ForwardsPassParallel (iLQR.cpp:824-934) batched over alphas: costs[B][nalpha] */
int tokp_forward_pass(tokp_traj* tr, const double* alphas, int nalpha, double* costs);
/* Optimiser::SaveBestRollout + iLQR::UpdateNominal (Optimiser.cpp:440-468, iLQR.cpp:936-948) */
int tokp_accept(tokp_traj* tr, int b, int alpha_idx);
/* iLQR::Optimise (iLQR.cpp:269-410) for all B trajectories in lock-step: U_out[B][T][m], stats[B][8] =
 * {initial_cost, final_cost, iterations, lambda, avg_pct_derivs, derivs_ms, bp_ms, fp_ms} (device-timed totals).
 * snapshot0 == U_init == NULL re-uses the device-resident inputs of the previous call; U_out / stats may be NULL. */
int tokp_optimise(tokp_traj* tr, const double* snapshot0, const double* U_init, int max_iterations, int min_iterations,
                  const double* lambda_init, double* U_out, double* stats);

/* named device-array readback (A, B, X_old, U_old, resid, r_x, r_u, l_x, l_xx, l_u, l_uu, K, k) for trajectory b;
 * returns count, -count if cap too small, -1e6 unknown */
int tokp_traj_get(tokp_traj* tr, int b, const char* name, double* out, int cap);
int tokp_traj_set(tokp_traj* tr, int b, const char* name, const double* in, int count);
/* switches of a trajectory batch: "fd_stage_sharing" (default 1: the control / velocity worlds of a key-point reuse the
 * nominal world's position / velocity stage like mj_stepSkip, Differentiator.cpp:118,155,253,276; 0: every world runs
 * the full step - same A / B bit for bit), "fd_lanes" (small hinge / slide models: 1 = those worlds run one
 * world per lane, 2 (default) = every FD world does and the group kernel only runs the position stage, 0 = off; same
 * A / B up to rounding), "fd_blocksync" (0..3, block-synchronous walk level of the FD kernels) */
int tokp_traj_set_option(tokp_traj* tr, const char* name, int value);
/* Per-trajectory step warnings, the analogue of mjData.warning: out[B][2]; out[b][0] = OR over all FD worlds of
 * trajectory b since the last clear, out[b][1] = the same over rollout worlds.  Bits: 1 bad qpos, 2 bad qvel, 4 bad qacc,
 * 8 constraint rows beyond the capacity (contacts dropped), 32 unsupported collision pair / contact dimension.
 * tokp_dynamics_derivs and tokp_optimise return TOKP_ERR_UNSUPPORTED / TOKP_ERR_STATE themselves when bit 32 (any
 * world) or bit 8 (FD worlds) is set. */
int tokp_traj_status(tokp_traj* tr, int* out, int clear);
/* counters of the last tokp_dynamics_derivs: [worlds, jacobians_x1e6, kernel_launches, worlds re-run with full row capacity] */
int tokp_traj_counters(tokp_traj* tr, int64_t out[4]);
/* CUDA-event time of the last derivative / backward / forward phase, ms (time_get_derivs_ms etc., iLQR.cpp:418-505) */
int tokp_timers(tokp_traj* tr, double ms[3]);
/* device time of the dominant kernel (k_fd_worlds, CUDA events on the launching stream):
 * out = {last launch ms, sum over launches ms, number of launches} */
int tokp_fd_kernel_ms(tokp_traj* tr, double out[3]);
/* live roofline denominators: out = {FP64 FMA TFLOP/s, device copy GB/s (read+write)} */
int tokp_measure_peaks(double out[2]);
/* total number of kernels this library launched since load */
int64_t tokp_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* TOKP_H_ */
