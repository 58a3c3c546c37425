// Compiled model ("model blob"): flat arrays mirroring the mjModel fields that MuJoCo's
// forward-dynamics pipeline reads for the model subset TrajOptKP's shipped tasks use
// (reference: MuJoCoHelper::InitSimulator loads the MJCF with mj_loadXML,
//  src/PhysicsSimulators/MuJoCoHelper.cpp:877-923).  Host-side only; the CUDA side consumes a
// device copy of the same arrays (see device_model.h).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace tokp {

enum JointType { JNT_FREE = 0, JNT_BALL = 1, JNT_SLIDE = 2, JNT_HINGE = 3 };
enum GeomType { GEOM_PLANE = 0, GEOM_HFIELD = 1, GEOM_SPHERE = 2, GEOM_CAPSULE = 3, GEOM_ELLIPSOID = 4,
                GEOM_CYLINDER = 5, GEOM_BOX = 6, GEOM_MESH = 7 };

// Field table: X(type, name, doc).  Every field is a std::vector<type>; int fields are int32,
// real fields are f64.  The same table drives serialisation, the C-ABI getters and the device upload.
#define TOKP_MODEL_INT_FIELDS(X)                                                                   \
  X(dims)               /* [nq,nv,nu,nbody,njnt,ngeom,npair,nmeshvert,maxcontact,maxefc,treedepth] */ \
  X(opt_int)            /* [iterations, ls_iterations, cone(0=pyramidal)] */                       \
  X(body_parentid) X(body_rootid) X(body_weldid) X(body_jntadr) X(body_jntnum)                     \
  X(body_dofadr) X(body_dofnum) X(body_geomadr) X(body_geomnum) X(body_depth)                      \
  X(body_subtreemask_lo) X(body_subtreemask_hi) /* bit j set: body j is in the subtree of i (incl. i) */ \
  X(jnt_type) X(jnt_qposadr) X(jnt_dofadr) X(jnt_bodyid) X(jnt_limited)                            \
  X(dof_bodyid) X(dof_jntid) X(dof_parentid)                                                       \
  X(dof_ancmask_lo) X(dof_ancmask_hi) /* bit j set: dof j is dof i or one of its ancestors */      \
  X(dof_treelast) /* last dof of the kinematic tree of dof i (trees are contiguous dof ranges) */   \
  X(geom_type) X(geom_bodyid) X(geom_contype) X(geom_conaffinity) X(geom_condim)                   \
  X(geom_priority) X(geom_meshadr) X(geom_meshnum)                                                 \
  X(pair_geom1) X(pair_geom2) X(pair_slotadr) /* first contact slot of the pair (prefix of its maximum contacts) */ \
  X(actuator_trnid) X(actuator_dofadr) X(actuator_ctrllimited) X(actuator_forcelimited)            \
  X(actuator_biastype)

#define TOKP_MODEL_REAL_FIELDS(X)                                                                  \
  X(opt_real)           /* [timestep, gx,gy,gz, tolerance, ls_tolerance, impratio, meaninertia] */ \
  X(qpos0)                                                                                         \
  X(body_pos) X(body_quat) X(body_ipos) X(body_iquat) X(body_mass) X(body_inertia)                 \
  X(body_invweight0) X(body_gravcomp)                                                              \
  X(jnt_pos) X(jnt_axis) X(jnt_range) X(jnt_solref) X(jnt_solimp) X(jnt_margin)                    \
  X(dof_armature) X(dof_damping) X(dof_invweight0)                                                 \
  X(geom_size) X(geom_pos) X(geom_quat) X(geom_friction) X(geom_solref) X(geom_solimp)             \
  X(geom_margin) X(geom_gap) X(geom_solmix) X(geom_rbound)                                         \
  X(mesh_vert)                                                                                     \
  X(actuator_gear) X(actuator_ctrlrange) X(actuator_forcerange) X(actuator_gainprm)                \
  X(actuator_biasprm)

struct Model {
#define X(name) std::vector<int32_t> name;
  TOKP_MODEL_INT_FIELDS(X)
#undef X
#define X(name) std::vector<double> name;
  TOKP_MODEL_REAL_FIELDS(X)
#undef X
  std::vector<std::string> body_names, jnt_names, geom_names, actuator_names, site_names;
  std::vector<int32_t> site_bodyid;
  std::vector<double> site_pos;

  int nq() const { return dims[0]; }
  int nv() const { return dims[1]; }
  int nu() const { return dims[2]; }
  int nbody() const { return dims[3]; }
  int njnt() const { return dims[4]; }
  int ngeom() const { return dims[5]; }
  int npair() const { return dims[6]; }
  int maxcontact() const { return dims[8]; }
  int maxefc() const { return dims[9]; }
  double timestep() const { return opt_real[0]; }

  // Serialise to a self-describing blob: "TOKPMDL1", then per field
  // [name: 32 bytes zero-padded][dtype: int32 (0=i32,1=f64)][count: int64][payload].
  std::vector<uint8_t> serialize() const;
};

// Compile an MJCF file (subset) into a Model.  `timestep` > 0 overrides <option timestep>
// exactly as MuJoCoHelper::InitSimulator does (MuJoCoHelper.cpp:895).  Throws std::runtime_error.
Model compile_mjcf_file(const std::string& path, double timestep);
Model compile_mjcf_string(const std::string& xml, const std::string& base_dir, double timestep);

}  // namespace tokp
