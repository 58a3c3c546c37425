// ORACLE (test infrastructure only - never linked into or called from the product path).
// Reader for the model blob produced by tokp::Model::serialize() (trajoptkp_b200/csrc/model.h).
// The blob is the input *data* (the equivalent of an mjModel); everything computed from it in
// oracle/ is an independent CPU restatement of MuJoCo's pipeline / the reference's algorithms.
#pragma once
#include <cstdint>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace ok {

enum { JNT_FREE = 0, JNT_BALL = 1, JNT_SLIDE = 2, JNT_HINGE = 3 };
enum { GEOM_PLANE = 0, GEOM_HFIELD = 1, GEOM_SPHERE = 2, GEOM_CAPSULE = 3, GEOM_ELLIPSOID = 4, GEOM_CYLINDER = 5,
       GEOM_BOX = 6, GEOM_MESH = 7 };

struct Model {
  std::map<std::string, std::vector<int32_t>> I;
  std::map<std::string, std::vector<double>> R;

  int nq, nv, nu, nbody, njnt, ngeom, npair, maxcontact, maxefc;
  double timestep, gravity[3], tolerance, ls_tolerance, impratio, meaninertia;
  int iterations, ls_iterations;

  // direct pointers for the hot loops
  const int32_t *body_parentid, *body_rootid, *body_weldid, *body_jntadr, *body_jntnum, *body_dofadr, *body_dofnum;
  const int32_t *jnt_type, *jnt_qposadr, *jnt_dofadr, *jnt_bodyid, *jnt_limited;
  const int32_t *dof_bodyid, *dof_jntid, *dof_parentid;
  const int32_t *geom_type, *geom_bodyid, *geom_condim, *geom_priority, *pair_geom1, *pair_geom2, *geom_meshadr, *geom_meshnum;
  const int32_t *actuator_dofadr, *actuator_ctrllimited, *actuator_forcelimited, *actuator_biastype, *actuator_trnid;
  const double *qpos0, *body_pos, *body_quat, *body_ipos, *body_iquat, *body_mass, *body_inertia, *body_invweight0,
      *body_gravcomp;
  const double *jnt_pos, *jnt_axis, *jnt_range, *jnt_solref, *jnt_solimp, *jnt_margin;
  const double *dof_armature, *dof_damping, *dof_invweight0;
  const double *geom_size, *geom_pos, *geom_quat, *geom_friction, *geom_solref, *geom_solimp, *geom_margin, *geom_gap,
      *geom_solmix, *geom_rbound, *mesh_vert;
  const double *actuator_gear, *actuator_ctrlrange, *actuator_forcerange, *actuator_gainprm, *actuator_biasprm;

  const int32_t* ip(const char* n) const {
    auto it = I.find(n);
    if (it == I.end()) throw std::runtime_error(std::string("oracle model: missing int field ") + n);
    return it->second.data();
  }
  const double* rp(const char* n) const {
    auto it = R.find(n);
    if (it == R.end()) throw std::runtime_error(std::string("oracle model: missing real field ") + n);
    return it->second.data();
  }

  void load(const uint8_t* blob, int64_t size) {
    if (size < 8 || std::memcmp(blob, "TOKPMDL1", 8) != 0) throw std::runtime_error("oracle model: bad magic");
    int64_t p = 8;
    while (p < size) {
      if (p + 44 > size) throw std::runtime_error("oracle model: truncated header");
      char nm[33];
      std::memcpy(nm, blob + p, 32);
      nm[32] = 0;
      int32_t dtype;
      int64_t cnt;
      std::memcpy(&dtype, blob + p + 32, 4);
      std::memcpy(&cnt, blob + p + 36, 8);
      p += 44;
      int64_t bytes = cnt * (dtype == 0 ? 4 : 8);
      if (p + bytes > size) throw std::runtime_error("oracle model: truncated payload");
      if (dtype == 0) {
        std::vector<int32_t> v(cnt);
        std::memcpy(v.data(), blob + p, bytes);
        I[nm] = std::move(v);
      } else {
        std::vector<double> v(cnt);
        std::memcpy(v.data(), blob + p, bytes);
        R[nm] = std::move(v);
      }
      p += bytes;
    }
    const int32_t* d = ip("dims");
    nq = d[0]; nv = d[1]; nu = d[2]; nbody = d[3]; njnt = d[4]; ngeom = d[5]; npair = d[6]; maxcontact = d[8]; maxefc = d[9];
    const int32_t* oi = ip("opt_int");
    iterations = oi[0]; ls_iterations = oi[1];
    const double* o = rp("opt_real");
    timestep = o[0]; gravity[0] = o[1]; gravity[1] = o[2]; gravity[2] = o[3];
    tolerance = o[4]; ls_tolerance = o[5]; impratio = o[6]; meaninertia = o[7];
#define P(n) n = ip(#n)
    P(body_parentid); P(body_rootid); P(body_weldid); P(body_jntadr); P(body_jntnum); P(body_dofadr); P(body_dofnum);
    P(jnt_type); P(jnt_qposadr); P(jnt_dofadr); P(jnt_bodyid); P(jnt_limited);
    P(dof_bodyid); P(dof_jntid); P(dof_parentid);
    P(geom_type); P(geom_bodyid); P(geom_condim); P(geom_priority); P(pair_geom1); P(pair_geom2); P(geom_meshadr); P(geom_meshnum);
    P(actuator_dofadr); P(actuator_ctrllimited); P(actuator_forcelimited); P(actuator_biastype); P(actuator_trnid);
#undef P
#define P(n) n = rp(#n)
    P(qpos0); P(body_pos); P(body_quat); P(body_ipos); P(body_iquat); P(body_mass); P(body_inertia); P(body_invweight0);
    P(body_gravcomp);
    P(jnt_pos); P(jnt_axis); P(jnt_range); P(jnt_solref); P(jnt_solimp); P(jnt_margin);
    P(dof_armature); P(dof_damping); P(dof_invweight0);
    P(geom_size); P(geom_pos); P(geom_quat); P(geom_friction); P(geom_solref); P(geom_solimp); P(geom_margin);
    P(geom_gap); P(geom_solmix); P(geom_rbound); P(mesh_vert);
    P(actuator_gear); P(actuator_ctrlrange); P(actuator_forcerange); P(actuator_gainprm); P(actuator_biasprm);
#undef P
  }
};

}  // namespace ok
