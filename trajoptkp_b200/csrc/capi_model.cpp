// C-ABI: model entry points (host only).  Replaces MuJoCoHelper::InitSimulator's mj_loadXML
// (src/PhysicsSimulators/MuJoCoHelper.cpp:877-923) and the mj_name2id lookups of MuJoCoHelper.cpp:36-438.
#include <cstring>
#include <exception>
#include <vector>

#include "host_math.h"
#include "tokp_internal.h"

namespace tokp {
static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
}  // namespace tokp

using tokp::set_error;

extern "C" {

const char* tokp_last_error(void) { return tokp::g_error.c_str(); }
const char* tokp_version(void) { return "tokp 0.1 (sm_100a)"; }

int tokp_model_compile(const char* mjcf_path, double timestep, tokp_model** out) {
  if (!mjcf_path || !out) { set_error("tokp_model_compile: null argument"); return TOKP_ERR_ARG; }
  try {
    auto* m = new tokp_model();
    m->host = tokp::compile_mjcf_file(mjcf_path, timestep);
    *out = m;
    return TOKP_OK;
  } catch (const std::exception& e) {
    set_error(e.what());
    return TOKP_ERR_MODEL;
  }
}

int tokp_model_compile_string(const char* mjcf_xml, const char* base_dir, double timestep, tokp_model** out) {
  if (!mjcf_xml || !out) { set_error("tokp_model_compile_string: null argument"); return TOKP_ERR_ARG; }
  try {
    auto* m = new tokp_model();
    m->host = tokp::compile_mjcf_string(mjcf_xml, base_dir ? base_dir : ".", timestep);
    *out = m;
    return TOKP_OK;
  } catch (const std::exception& e) {
    set_error(e.what());
    return TOKP_ERR_MODEL;
  }
}

void tokp_model_destroy(tokp_model* m) {
  if (!m) return;
  if (m->dev) tokp::device_model_free(m->dev);
  delete m;
}

int tokp_model_dims(const tokp_model* m, int dims[11]) {
  if (!m || !dims) { set_error("tokp_model_dims: null argument"); return TOKP_ERR_ARG; }
  for (int i = 0; i < 11; i++) dims[i] = m->host.dims[i];
  return TOKP_OK;
}

int tokp_model_name2id(const tokp_model* m, int objtype, const char* name) {
  if (!m || !name) return -1;
  const std::vector<std::string>* names = nullptr;
  switch (objtype) {
    case TOKP_OBJ_BODY: names = &m->host.body_names; break;
    case TOKP_OBJ_JOINT: names = &m->host.jnt_names; break;
    case TOKP_OBJ_GEOM: names = &m->host.geom_names; break;
    case TOKP_OBJ_SITE: names = &m->host.site_names; break;
    case TOKP_OBJ_ACTUATOR: names = &m->host.actuator_names; break;
    default: return -1;
  }
  for (size_t i = 0; i < names->size(); i++)
    if ((*names)[i] == name) return (int)i;
  return -1;
}

int64_t tokp_model_blob(const tokp_model* m, void* buf, int64_t cap) {
  if (!m) return -1;
  std::vector<uint8_t> blob = m->host.serialize();
  if (buf) {
    if ((int64_t)blob.size() > cap) return -1;
    std::memcpy(buf, blob.data(), blob.size());
  }
  return (int64_t)blob.size();
}

int tokp_model_get_real(const tokp_model* m, const char* field, double* out, int cap) {
  if (!m || !field) return -1000000;
  std::string nm(field);
#define X(name)                                                        \
  if (nm == #name) {                                                   \
    int c = (int)m->host.name.size();                                  \
    if (c > cap || !out) return -c;                                    \
    std::memcpy(out, m->host.name.data(), sizeof(double) * c);         \
    return c;                                                          \
  }
  TOKP_MODEL_REAL_FIELDS(X)
#undef X
  return -1000000;
}

int tokp_model_get_int(const tokp_model* m, const char* field, int32_t* out, int cap) {
  if (!m || !field) return -1000000;
  std::string nm(field);
#define X(name)                                                        \
  if (nm == #name) {                                                   \
    int c = (int)m->host.name.size();                                  \
    if (c > cap || !out) return -c;                                    \
    std::memcpy(out, m->host.name.data(), sizeof(int32_t) * c);        \
    return c;                                                          \
  }
  TOKP_MODEL_INT_FIELDS(X)
#undef X
  return -1000000;
}

int tokp_model_body_pose(const tokp_model* m, const double* qpos, int body, double pos[3], double quat[4]) {
  if (!m || !qpos || !pos || !quat || body < 0 || body >= m->host.nbody()) { set_error("tokp_model_body_pose: bad argument"); return TOKP_ERR_ARG; }
  namespace hm = tokp::hm;
  const auto& h = m->host;
  std::vector<int> chain;
  for (int b = body; b > 0; b = h.body_parentid[b]) chain.push_back(b);
  double xpos[3] = {0, 0, 0}, xquat[4] = {1, 0, 0, 0};
  for (int c = (int)chain.size() - 1; c >= 0; c--) {
    int b = chain[c];
    int jadr = h.body_jntadr[b], jnum = h.body_jntnum[b];
    if (jnum == 1 && h.jnt_type[jadr] == tokp::JNT_FREE) {
      int qa = h.jnt_qposadr[jadr];
      for (int k = 0; k < 3; k++) xpos[k] = qpos[qa + k];
      for (int k = 0; k < 4; k++) xquat[k] = qpos[qa + 3 + k];
      hm::quat_normalize(xquat);
      continue;
    }
    double p[3], q[4];
    hm::rot_vec_quat(p, &h.body_pos[3 * b], xquat);
    for (int k = 0; k < 3; k++) p[k] += xpos[k];
    hm::quat_mul(q, xquat, &h.body_quat[4 * b]);
    for (int j = jadr; j >= 0 && j < jadr + jnum; j++) {
      int qa = h.jnt_qposadr[j];
      double axis[3], anchor[3];
      hm::rot_vec_quat(axis, &h.jnt_axis[3 * j], q);
      hm::rot_vec_quat(anchor, &h.jnt_pos[3 * j], q);
      for (int k = 0; k < 3; k++) anchor[k] += p[k];
      if (h.jnt_type[j] == tokp::JNT_SLIDE) {
        double dq = qpos[qa] - h.qpos0[qa];
        for (int k = 0; k < 3; k++) p[k] += axis[k] * dq;
      } else if (h.jnt_type[j] == tokp::JNT_HINGE) {
        double ql[4], t[4], v[3];
        hm::axis_angle2quat(ql, &h.jnt_axis[3 * j], qpos[qa] - h.qpos0[qa]);
        hm::quat_mul(t, q, ql);
        for (int k = 0; k < 4; k++) q[k] = t[k];
        hm::rot_vec_quat(v, &h.jnt_pos[3 * j], q);
        for (int k = 0; k < 3; k++) p[k] = anchor[k] - v[k];
      } else {
        set_error("tokp_model_body_pose: unsupported joint type on the chain");
        return TOKP_ERR_UNSUPPORTED;
      }
    }
    hm::quat_normalize(q);
    for (int k = 0; k < 3; k++) xpos[k] = p[k];
    for (int k = 0; k < 4; k++) xquat[k] = q[k];
  }
  for (int k = 0; k < 3; k++) pos[k] = xpos[k];
  for (int k = 0; k < 4; k++) quat[k] = xquat[k];
  return TOKP_OK;
}

int tokp_snapshot_size(const tokp_model* m) {
  if (!m) return -1;
  const auto& h = m->host;
  return 1 + h.nq() + 4 * h.nv() + 6 * h.nbody() + h.nu();
}

}  // extern "C"
