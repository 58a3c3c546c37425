// Device-resident copy of the compiled model: plain pointers into one HBM allocation, passed to
// kernels by value.  Field meaning = mjModel's (see model.h).
#pragma once
#include <cstdint>

struct DM {
  int nq, nv, nu, nbody, njnt, ngeom, npair, maxcontact, maxefc, treedepth;
  int iterations, ls_iterations;
  int any_damping, any_gravcomp;
  int spec_nv;    // nv when the model may run a kernel instantiation specialised on nv (those carry no cylinder / mesh /
                  // convex-collider code: every collision pair must be plane-{sphere, capsule, box}); 0 = generic kernels
  int lane_ok;    // the lane-per-world FD kernel applies: spec_nv in {2, 9}, hinge / slide joints only, no gravity
                  // compensation, at most 64 constraint rows
  // tuning overrides, read from the environment ONCE (model upload; again at tokp_traj_create for the batch's own copy):
  // TOKP_WARPS (warps per block of the world kernels, 0 = planned), TOKP_EFC_CAP (fast constraint-row capacity, 0 = 32)
  int opt_warps, opt_efc_cap;
  int fast_ecap;  // constraint-row capacity of the fast pass (0 = default 32); a trajectory batch raises its own copy
                  // when many of its FD worlds overflow (update_fd_counters)
  double timestep, gravity[3], tolerance, ls_tolerance, meaninertia;
  // int arrays
  const int *body_parentid, *body_rootid, *body_weldid, *body_jntadr, *body_jntnum, *body_dofadr, *body_dofnum, *body_depth;
  const unsigned *body_subtreemask_lo, *body_subtreemask_hi;
  const int *jnt_type, *jnt_qposadr, *jnt_dofadr, *jnt_bodyid, *jnt_limited;
  const int *dof_bodyid, *dof_jntid, *dof_parentid;
  const unsigned *dof_ancmask_lo, *dof_ancmask_hi;
  const int* dof_treelast;  // last dof of the dof's kinematic tree
  int max_tree_dofs;        // dofs of the largest kinematic tree (= nv for a single tree)
  const int *geom_type, *geom_bodyid, *geom_condim, *geom_priority, *pair_geom1, *pair_geom2, *pair_slotadr, *geom_meshadr, *geom_meshnum;
  const int *actuator_trnid, *actuator_dofadr, *actuator_ctrllimited, *actuator_forcelimited, *actuator_biastype;
  // real arrays
  const double *qpos0, *body_pos, *body_quat, *body_ipos, *body_iquat, *body_mass, *body_inertia, *body_invweight0,
      *body_gravcomp;
  const double *jnt_pos, *jnt_axis, *jnt_range, *jnt_solref, *jnt_solimp, *jnt_margin;
  const double *dof_armature, *dof_damping, *dof_invweight0;
  const double *geom_size, *geom_pos, *geom_quat, *geom_friction, *geom_solref, *geom_solimp, *geom_margin, *geom_gap,
      *geom_solmix, *geom_rbound, *mesh_vert;
  const double *actuator_gear, *actuator_ctrlrange, *actuator_forcerange, *actuator_gainprm, *actuator_biasprm;
};

struct DeviceModel {
  DM dm;
  void* arena = nullptr;  // single device allocation holding every array
  int device = -1;        // CUDA device the arena lives on: the handle is valid on that device only
  int lanes = 0;          // lanes per world of tokp_step_worlds (TOKP_LANES read once, on the first call)
};
