// MJCF-subset compiler: XML -> tokp::Model.
//
// Reproduces the parts of MuJoCo's model compiler the shipped TrajOptKP tasks rely on
// (reference: MuJoCoHelper::InitSimulator -> mj_loadXML, src/PhysicsSimulators/MuJoCoHelper.cpp:877-923;
//  models listed in SURVEY.md Appendix B): default classes + childclass, compiler angle/eulerseq,
// body tree, hinge/slide/free joints, plane/sphere/capsule/cylinder/box geoms (fromto, zaxis, euler, quat),
// inertia-from-geoms, motors / position actuators, joint limits, per-geom contact parameters, and the
// compile-time constants MuJoCo derives at qpos0 (dof_invweight0, body_invweight0, stat.meaninertia).
#include <algorithm>
#include <array>
#include <cctype>
#include <cmath>
#include <cstring>
#include <fstream>
#include <functional>
#include <map>
#include <sstream>
#include <stdexcept>

#include "host_math.h"
#include "model.h"
#include "xml_min.h"

namespace tokp {
namespace {

constexpr double kPi = 3.14159265358979323846;
constexpr double kMinVal = 1e-15;

using AttrMap = std::map<std::string, std::string>;

struct DefaultClass {
  std::string name;
  std::map<std::string, AttrMap> elems;  // "joint","geom","site","actuator","mesh" -> attrs
};

std::vector<double> parse_reals(const std::string& s) {
  std::vector<double> out;
  std::istringstream is(s);
  double v;
  while (is >> v) out.push_back(v);
  return out;
}

bool parse_bool(const std::string& s) { return s == "true" || s == "1"; }

struct Compiler {
  Model m;
  std::map<std::string, DefaultClass> classes;
  bool angle_degree = true;
  std::string eulerseq = "xyz";
  bool autolimits = true;
  std::string base_dir;
  double opt_timestep = 0.002;
  double gravity[3] = {0, 0, -9.81};
  int opt_iterations = 100, opt_ls_iterations = 50;
  double opt_tolerance = 1e-8, opt_ls_tolerance = 0.01, opt_impratio = 1.0;

  // per-geom data gathered during parsing
  struct GeomTmp { double mass; double inertia[3]; };
  std::vector<GeomTmp> geom_tmp;
  struct InertialTmp { bool given = false; double pos[3], quat[4], mass, inertia[3]; };
  std::vector<InertialTmp> inertial_tmp;

  // ---- defaults ---------------------------------------------------------------------------
  static std::string default_key(const std::string& tag) {
    if (tag == "motor" || tag == "position" || tag == "velocity" || tag == "general") return "actuator";
    return tag;
  }

  void parse_default(const XmlNode& node, const DefaultClass* parent, bool top) {
    DefaultClass cls;
    if (parent) cls = *parent;
    else if (top && classes.count("main")) cls = classes["main"];  // several top-level <default> blocks merge
    const std::string* nm = node.attr("class");
    cls.name = (top || !nm) ? (nm ? *nm : "main") : *nm;
    if (top && !nm) cls.name = "main";
    for (auto& ch : node.children) {
      if (ch->tag == "default") continue;
      AttrMap& dst = cls.elems[default_key(ch->tag)];
      if (ch->tag == "motor") {  // motor shortcut: fixed gain 1, no bias
        dst["__kind"] = "motor";
      } else if (ch->tag == "position") {
        dst["__kind"] = "position";
      }
      for (auto& a : ch->attrs) dst[a.first] = a.second;
    }
    classes[cls.name] = cls;
    for (auto& ch : node.children)
      if (ch->tag == "default") parse_default(*ch, &classes[cls.name], false);
  }

  // attribute lookup: element attr, else class default
  struct Getter {
    const XmlNode* node;
    const AttrMap* defs;
    const std::string* get(const std::string& k) const {
      if (auto* a = node->attr(k)) return a;
      if (defs) {
        auto it = defs->find(k);
        if (it != defs->end()) return &it->second;
      }
      return nullptr;
    }
    bool has(const std::string& k) const { return get(k) != nullptr; }
    std::vector<double> reals(const std::string& k) const {
      auto* a = get(k);
      return a ? parse_reals(*a) : std::vector<double>();
    }
    double real(const std::string& k, double def) const {
      auto v = reals(k);
      return v.empty() ? def : v[0];
    }
    int integer(const std::string& k, int def) const {
      auto v = reals(k);
      return v.empty() ? def : (int)std::lround(v[0]);
    }
    std::string str(const std::string& k, const std::string& def) const {
      auto* a = get(k);
      return a ? *a : def;
    }
  };

  Getter getter(const XmlNode& node, const std::string& tag, const std::string& childclass) {
    std::string cname = "main";
    if (auto* c = node.attr("class")) cname = *c;
    else if (!childclass.empty()) cname = childclass;
    auto it = classes.find(cname);
    if (it == classes.end()) {
      if (cname != "main") throw std::runtime_error("mjcf: unknown default class '" + cname + "'");
      return Getter{&node, nullptr};
    }
    auto e = it->second.elems.find(default_key(tag));
    return Getter{&node, e == it->second.elems.end() ? nullptr : &e->second};
  }

  // ---- mesh assets (MuJoCo: mjCMesh::LoadSTL / LoadOBJ / Process) -----------------------------
  // A mesh is loaded lazily, the first time a geom that collides or carries mass refers to it.  Processing follows
  // MuJoCo's compiler: repeated vertices are merged, volume / centre of mass / inertia come from the faces
  // (tetrahedra about the vertex centroid, absolute volumes), and the vertices are re-expressed in the frame that
  // is centred at the centre of mass and aligned with the principal axes; the geom frame absorbs that transform.
  // Collision uses the convex hull, which a support function over all vertices reproduces, so no hull is built.
  struct MeshAsset {
    std::string file;
    double scale[3] = {1, 1, 1};
    bool loaded = false;
    std::vector<double> vert;  // processed vertices (mesh frame), 3 per vertex
    double pos[3] = {0, 0, 0}, quat[4] = {1, 0, 0, 0};  // mesh frame in the file's coordinates
    double volume = 0, inertia_unit[3] = {0, 0, 0};      // principal inertia at unit density
    double half[3] = {0, 0, 0};                          // half-extents of the axis-aligned box in the mesh frame
    int vertadr = -1;                                    // offset (vertices) into Model::mesh_vert once used
  };
  std::map<std::string, MeshAsset> meshes;
  std::string meshdir;

  void parse_assets(const XmlNode& node) {
    for (auto& ch : node.children) {
      if (ch->tag != "mesh") continue;
      const std::string* file = ch->attr("file");
      if (!file) throw std::runtime_error("mjcf: <mesh> without file (inline vertex data is outside the subset)");
      std::string name;
      if (auto* n = ch->attr("name")) name = *n;
      else {
        size_t s0 = file->find_last_of('/'), dot = file->find_last_of('.');
        s0 = s0 == std::string::npos ? 0 : s0 + 1;
        name = file->substr(s0, dot == std::string::npos || dot < s0 ? std::string::npos : dot - s0);
      }
      MeshAsset a;
      a.file = *file;
      if (auto* sc = ch->attr("scale")) {
        auto v = parse_reals(*sc);
        if (v.size() == 3) for (int i = 0; i < 3; i++) a.scale[i] = v[i];
      }
      meshes[name] = a;
    }
  }

  static void read_stl(const std::string& path, std::vector<float>& tri) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("mjcf: cannot open mesh '" + path + "'");
    std::vector<char> buf((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    if (buf.size() < 84) throw std::runtime_error("mjcf: '" + path + "' is not a binary STL");
    uint32_t n;
    std::memcpy(&n, buf.data() + 80, 4);
    if (buf.size() < 84 + (size_t)n * 50) throw std::runtime_error("mjcf: '" + path + "' is not a binary STL (ASCII STL unsupported)");
    tri.resize((size_t)n * 9);
    for (uint32_t t = 0; t < n; t++) std::memcpy(&tri[(size_t)t * 9], buf.data() + 84 + (size_t)t * 50 + 12, 36);
  }

  static void read_obj(const std::string& path, std::vector<double>& v, std::vector<int>& face) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("mjcf: cannot open mesh '" + path + "'");
    std::string line;
    while (std::getline(f, line)) {
      std::istringstream is(line);
      std::string key;
      is >> key;
      if (key == "v") {
        double x, y, z;
        if (is >> x >> y >> z) { v.push_back(x); v.push_back(y); v.push_back(z); }
      } else if (key == "f") {
        std::vector<int> idx;
        std::string tok;
        while (is >> tok) {
          int i = std::atoi(tok.c_str());  // "a/b/c": atoi stops at '/'
          int nvert = (int)v.size() / 3;
          idx.push_back(i > 0 ? i - 1 : nvert + i);
        }
        for (size_t k = 2; k < idx.size(); k++) { face.push_back(idx[0]); face.push_back(idx[k - 1]); face.push_back(idx[k]); }
      }
    }
  }

  MeshAsset& load_mesh(const std::string& name) {
    auto it = meshes.find(name);
    if (it == meshes.end()) throw std::runtime_error("mjcf: unknown mesh '" + name + "'");
    MeshAsset& a = it->second;
    if (a.loaded) return a;
    std::string path = base_dir + "/" + (meshdir.empty() ? "" : meshdir + "/") + a.file;
    std::vector<double> v;
    std::vector<int> face;
    std::string ext = a.file.size() >= 4 ? a.file.substr(a.file.size() - 4) : "";
    for (auto& c : ext) c = (char)std::tolower(c);
    if (ext == ".stl") {
      std::vector<float> tri;
      read_stl(path, tri);
      // merge repeated vertices (exact float equality, as written by the exporter)
      std::map<std::array<float, 3>, int> seen;
      for (size_t k = 0; k < tri.size() / 3; k++) {
        std::array<float, 3> key = {tri[3 * k], tri[3 * k + 1], tri[3 * k + 2]};
        auto ins = seen.emplace(key, (int)v.size() / 3);
        if (ins.second) { v.push_back(key[0]); v.push_back(key[1]); v.push_back(key[2]); }
        face.push_back(ins.first->second);
      }
    } else if (ext == ".obj") {
      read_obj(path, v, face);
      for (int idx : face)
        if (idx < 0 || idx >= (int)v.size() / 3)
          throw std::runtime_error("mjcf: mesh '" + name + "' has a face index outside its vertex list");
    } else {
      throw std::runtime_error("mjcf: mesh format of '" + a.file + "' unsupported (binary STL and OBJ only)");
    }
    int nvert = (int)v.size() / 3, nface = (int)face.size() / 3;
    if (nvert < 4 || nface < 4) throw std::runtime_error("mjcf: mesh '" + name + "' is degenerate");
    for (int i = 0; i < nvert; i++) for (int k = 0; k < 3; k++) v[3 * i + k] *= a.scale[k];
    // volume and centre of mass
    double c0[3] = {0, 0, 0};
    for (int i = 0; i < nvert; i++) for (int k = 0; k < 3; k++) c0[k] += v[3 * i + k] / nvert;
    double vol = 0, com[3] = {0, 0, 0};
    auto tet = [&](int f, const double* o, double p[3][3]) {
      for (int j = 0; j < 3; j++) for (int k = 0; k < 3; k++) p[j][k] = v[3 * face[3 * f + j] + k] - o[k];
      double cr[3];
      hm::cross(cr, p[1], p[2]);
      return std::fabs(hm::dot3(p[0], cr)) / 6.0;
    };
    for (int f = 0; f < nface; f++) {
      double p[3][3];
      double tv = tet(f, c0, p);
      vol += tv;
      for (int k = 0; k < 3; k++) com[k] += tv * (p[0][k] + p[1][k] + p[2][k]) / 4.0;
    }
    if (vol < kMinVal) throw std::runtime_error("mjcf: mesh '" + name + "' has no volume");
    for (int k = 0; k < 3; k++) com[k] = c0[k] + com[k] / vol;
    // second moments about the centre of mass: per tetrahedron (apex c0) V/20 (sum p p^T + s s^T)
    double C[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int f = 0; f < nface; f++) {
      double p[3][3];
      double tv = tet(f, c0, p);
      double q[4][3], sum[3] = {0, 0, 0};
      for (int k = 0; k < 3; k++) q[3][k] = c0[k] - com[k];
      for (int j = 0; j < 3; j++) for (int k = 0; k < 3; k++) q[j][k] = p[j][k] + c0[k] - com[k];
      for (int j = 0; j < 4; j++) for (int k = 0; k < 3; k++) sum[k] += q[j][k];
      for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) {
          double acc = sum[r] * sum[c];
          for (int j = 0; j < 4; j++) acc += q[j][r] * q[j][c];
          C[3 * r + c] += tv / 20.0 * acc;
        }
    }
    double trC = C[0] + C[4] + C[8];
    double I[9];
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) I[3 * r + c] = (r == c ? trC : 0.0) - C[3 * r + c];
    double ev[3], R[9];
    hm::eig3(I, ev, R);
    a.volume = vol;
    for (int k = 0; k < 3; k++) { a.inertia_unit[k] = ev[k]; a.pos[k] = com[k]; }
    hm::mat2quat(a.quat, R);
    a.vert.resize(v.size());
    for (int i = 0; i < nvert; i++) {
      double d[3] = {v[3 * i] - com[0], v[3 * i + 1] - com[1], v[3 * i + 2] - com[2]};
      for (int k = 0; k < 3; k++) {  // mesh frame: columns of R are the principal axes
        double x = R[k] * d[0] + R[3 + k] * d[1] + R[6 + k] * d[2];
        a.vert[3 * i + k] = x;
        a.half[k] = std::max(a.half[k], std::fabs(x));
      }
    }
    a.loaded = true;
    return a;
  }

  // ---- orientation ------------------------------------------------------------------------
  double to_rad(double a) const { return angle_degree ? a * kPi / 180.0 : a; }

  // only attributes given on the element itself define orientation (MuJoCo applies defaults for
  // orientation too, but no shipped model uses that)
  void orientation(const Getter& g, double quat[4]) {
    quat[0] = 1; quat[1] = quat[2] = quat[3] = 0;
    if (g.has("quat")) {
      auto v = g.reals("quat");
      if (v.size() != 4) throw std::runtime_error("mjcf: quat needs 4 numbers");
      for (int i = 0; i < 4; i++) quat[i] = v[i];
      hm::quat_normalize(quat);
    } else if (g.has("euler")) {
      auto v = g.reals("euler");
      if (v.size() != 3) throw std::runtime_error("mjcf: euler needs 3 numbers");
      // MuJoCo: lower-case sequence = intrinsic rotations, composed left to right
      for (int i = 0; i < 3; i++) {
        double axis[3] = {0, 0, 0};
        char c = eulerseq[i];
        int ax = (c == 'x' || c == 'X') ? 0 : (c == 'y' || c == 'Y') ? 1 : 2;
        axis[ax] = 1;
        double qi[4], tmp[4];
        hm::axis_angle2quat(qi, axis, to_rad(v[i]));
        if (c >= 'a') hm::quat_mul(tmp, quat, qi);  // intrinsic
        else hm::quat_mul(tmp, qi, quat);           // extrinsic
        std::memcpy(quat, tmp, sizeof(tmp));
      }
      hm::quat_normalize(quat);
    } else if (g.has("axisangle")) {
      auto v = g.reals("axisangle");
      if (v.size() != 4) throw std::runtime_error("mjcf: axisangle needs 4 numbers");
      double axis[3] = {v[0], v[1], v[2]};
      hm::normalize3(axis);
      hm::axis_angle2quat(quat, axis, to_rad(v[3]));
    } else if (g.has("zaxis")) {
      auto v = g.reals("zaxis");
      if (v.size() != 3) throw std::runtime_error("mjcf: zaxis needs 3 numbers");
      hm::z2quat(quat, v.data());
    } else if (g.has("xyaxes")) {
      auto v = g.reals("xyaxes");
      if (v.size() != 6) throw std::runtime_error("mjcf: xyaxes needs 6 numbers");
      double x[3] = {v[0], v[1], v[2]}, y[3] = {v[3], v[4], v[5]}, z[3];
      hm::normalize3(x);
      double d = hm::dot3(x, y);
      for (int i = 0; i < 3; i++) y[i] -= d * x[i];
      hm::normalize3(y);
      hm::cross(z, x, y);
      double R[9] = {x[0], y[0], z[0], x[1], y[1], z[1], x[2], y[2], z[2]};
      hm::mat2quat(quat, R);
    }
  }

  // ---- element parsers ----------------------------------------------------------------------
  template <class T>
  static void push_n(std::vector<T>& v, const T* p, int n) { v.insert(v.end(), p, p + n); }

  void add_joint(const XmlNode& node, int body, const std::string& childclass, bool freejoint) {
    Getter g = getter(node, "joint", childclass);
    if (freejoint) g.defs = nullptr;
    std::string type = freejoint ? "free" : g.str("type", "hinge");
    int jt = type == "free" ? JNT_FREE : type == "ball" ? JNT_BALL : type == "slide" ? JNT_SLIDE : JNT_HINGE;
    if (jt == JNT_BALL) throw std::runtime_error("mjcf: ball joints are outside the supported subset");
    int jid = (int)m.jnt_type.size();
    m.jnt_type.push_back(jt);
    m.jnt_bodyid.push_back(body);
    m.jnt_names.push_back(g.str("name", ""));
    int nq = (int)m.qpos0.size(), nv = (int)m.dof_bodyid.size();
    m.jnt_qposadr.push_back(nq);
    m.jnt_dofadr.push_back(nv);
    double pos[3] = {0, 0, 0}, axis[3] = {0, 0, 1};
    if (jt != JNT_FREE) {
      auto p = g.reals("pos");
      if (p.size() == 3) for (int i = 0; i < 3; i++) pos[i] = p[i];
      auto a = g.reals("axis");
      if (a.size() == 3) for (int i = 0; i < 3; i++) axis[i] = a[i];
      hm::normalize3(axis);
    }
    push_n(m.jnt_pos, pos, 3);
    push_n(m.jnt_axis, axis, 3);
    // limits
    double range[2] = {0, 0};
    bool has_range = false;
    auto r = g.reals("range");
    if (jt != JNT_FREE && r.size() == 2) {
      has_range = true;
      range[0] = r[0]; range[1] = r[1];
      if (jt == JNT_HINGE) { range[0] = to_rad(range[0]); range[1] = to_rad(range[1]); }
    }
    std::string lim = g.str("limited", "auto");
    int limited = (lim == "auto") ? (autolimits && has_range) : parse_bool(lim);
    if (jt == JNT_FREE) limited = 0;
    m.jnt_limited.push_back(limited);
    push_n(m.jnt_range, range, 2);
    double solref[2] = {0.02, 1}, solimp[5] = {0.9, 0.95, 0.001, 0.5, 2};
    auto sr = g.reals("solreflimit");
    for (size_t i = 0; i < sr.size() && i < 2; i++) solref[i] = sr[i];
    auto si = g.reals("solimplimit");
    for (size_t i = 0; i < si.size() && i < 5; i++) solimp[i] = si[i];
    push_n(m.jnt_solref, solref, 2);
    push_n(m.jnt_solimp, solimp, 5);
    m.jnt_margin.push_back(g.real("margin", 0.0));
    // dofs / qpos
    if (g.real("frictionloss", 0.0) != 0.0 || g.real("stiffness", 0.0) != 0.0)
      throw std::runtime_error("mjcf: joint frictionloss / stiffness are outside the supported subset");
    double armature = jt == JNT_FREE ? 0.0 : g.real("armature", 0.0);
    double damping = jt == JNT_FREE ? 0.0 : g.real("damping", 0.0);
    int ndof = jt == JNT_FREE ? 6 : 1;
    for (int k = 0; k < ndof; k++) {
      m.dof_bodyid.push_back(body);
      m.dof_jntid.push_back(jid);
      m.dof_armature.push_back(armature);
      m.dof_damping.push_back(damping);
    }
    if (jt == JNT_FREE) {
      // qpos0 = body pos/quat in the parent (world) frame
      for (int k = 0; k < 3; k++) m.qpos0.push_back(m.body_pos[3 * body + k]);
      for (int k = 0; k < 4; k++) m.qpos0.push_back(m.body_quat[4 * body + k]);
    } else {
      double ref = g.real("ref", 0.0);
      if (jt == JNT_HINGE) ref = to_rad(ref);
      m.qpos0.push_back(ref);
    }
  }

  void add_geom(const XmlNode& node, int body, const std::string& childclass) {
    Getter g = getter(node, "geom", childclass);
    std::string type = g.str("type", "sphere");
    int gt;
    if (type == "plane") gt = GEOM_PLANE;
    else if (type == "sphere") gt = GEOM_SPHERE;
    else if (type == "capsule") gt = GEOM_CAPSULE;
    else if (type == "cylinder") gt = GEOM_CYLINDER;
    else if (type == "box") gt = GEOM_BOX;
    else if (type == "mesh") gt = GEOM_MESH;
    else throw std::runtime_error("mjcf: geom type '" + type + "' is outside the supported subset");
    double size[3] = {0, 0, 0};
    auto sz = g.reals("size");
    for (size_t i = 0; i < sz.size() && i < 3; i++) size[i] = sz[i];
    double pos[3] = {0, 0, 0}, quat[4] = {1, 0, 0, 0};
    auto p = Getter{&node, nullptr}.reals("pos");
    if (p.size() == 3) for (int i = 0; i < 3; i++) pos[i] = p[i];
    orientation(Getter{&node, nullptr}, quat);
    if (node.has("fromto")) {
      auto ft = parse_reals(*node.attr("fromto"));
      if (ft.size() != 6) throw std::runtime_error("mjcf: fromto needs 6 numbers");
      double vec[3] = {ft[0] - ft[3], ft[1] - ft[4], ft[2] - ft[5]};
      double len = hm::norm3(vec);
      for (int i = 0; i < 3; i++) pos[i] = 0.5 * (ft[i] + ft[i + 3]);
      hm::z2quat(quat, vec);
      if (gt == GEOM_BOX) size[2] = 0.5 * len; else size[1] = 0.5 * len;
    }
    // mesh geoms: visual-only ones (no collision bits, zero mass) keep their slot but never open the file
    MeshAsset* mesh = nullptr;
    int contype = g.integer("contype", 1), conaffinity = g.integer("conaffinity", 1);
    if (gt == GEOM_MESH) {
      bool massless = g.has("mass") && g.real("mass", 0) == 0.0;
      if (contype != 0 || conaffinity != 0 || !massless) {
        mesh = &load_mesh(g.str("mesh", ""));
        if (mesh->vertadr < 0) {
          mesh->vertadr = (int)m.mesh_vert.size() / 3;
          m.mesh_vert.insert(m.mesh_vert.end(), mesh->vert.begin(), mesh->vert.end());
        }
        // the geom frame absorbs the mesh frame (centre of mass, principal axes)
        double off[3], q2[4];
        hm::rot_vec_quat(off, mesh->pos, quat);
        for (int i = 0; i < 3; i++) pos[i] += off[i];
        hm::quat_mul(q2, quat, mesh->quat);
        std::memcpy(quat, q2, sizeof(q2));
        for (int i = 0; i < 3; i++) size[i] = mesh->half[i];
      }
    }
    m.geom_type.push_back(gt);
    m.geom_bodyid.push_back(body);
    m.geom_names.push_back(g.str("name", ""));
    push_n(m.geom_size, size, 3);
    push_n(m.geom_pos, pos, 3);
    push_n(m.geom_quat, quat, 4);
    m.geom_contype.push_back(contype);
    m.geom_conaffinity.push_back(conaffinity);
    m.geom_condim.push_back(g.integer("condim", 3));
    m.geom_priority.push_back(g.integer("priority", 0));
    m.geom_meshadr.push_back(mesh ? mesh->vertadr : -1);
    m.geom_meshnum.push_back(mesh ? (int)mesh->vert.size() / 3 : 0);
    double friction[3] = {1, 0.005, 0.0001};
    auto fr = g.reals("friction");
    for (size_t i = 0; i < fr.size() && i < 3; i++) friction[i] = fr[i];
    push_n(m.geom_friction, friction, 3);
    double solref[2] = {0.02, 1}, solimp[5] = {0.9, 0.95, 0.001, 0.5, 2};
    auto sr = g.reals("solref");
    for (size_t i = 0; i < sr.size() && i < 2; i++) solref[i] = sr[i];
    auto si = g.reals("solimp");
    for (size_t i = 0; i < si.size() && i < 5; i++) solimp[i] = si[i];
    push_n(m.geom_solref, solref, 2);
    push_n(m.geom_solimp, solimp, 5);
    m.geom_margin.push_back(g.real("margin", 0.0));
    m.geom_gap.push_back(g.real("gap", 0.0));
    m.geom_solmix.push_back(g.real("solmix", 1.0));
    // bounding sphere radius
    double rb = 0;
    switch (gt) {
      case GEOM_SPHERE: rb = size[0]; break;
      case GEOM_CAPSULE: rb = size[0] + size[1]; break;
      case GEOM_CYLINDER: rb = std::sqrt(size[0] * size[0] + size[1] * size[1]); break;
      case GEOM_BOX: rb = std::sqrt(size[0] * size[0] + size[1] * size[1] + size[2] * size[2]); break;
      case GEOM_MESH: rb = std::sqrt(size[0] * size[0] + size[1] * size[1] + size[2] * size[2]); break;  // box bound
      default: rb = 0;
    }
    m.geom_rbound.push_back(rb);
    // mass & inertia about the geom's own frame (MuJoCo: mjCGeom::GetVolume / SetInertia)
    GeomTmp t{0, {0, 0, 0}};
    double vol = 0;
    switch (gt) {
      case GEOM_SPHERE: vol = 4.0 * kPi * size[0] * size[0] * size[0] / 3.0; break;
      case GEOM_CAPSULE: {
        double h = 2 * size[1];
        vol = kPi * (size[0] * size[0] * h + 4.0 * size[0] * size[0] * size[0] / 3.0);
        break;
      }
      case GEOM_CYLINDER: vol = kPi * size[0] * size[0] * 2 * size[1]; break;
      case GEOM_BOX: vol = 8 * size[0] * size[1] * size[2]; break;
      case GEOM_MESH: vol = mesh ? mesh->volume : 0; break;
      default: vol = 0;
    }
    if (g.has("mass")) t.mass = g.real("mass", 0);
    else t.mass = g.real("density", 1000.0) * vol;
    if (gt == GEOM_PLANE) t.mass = 0;
    double mass = t.mass;
    switch (gt) {
      case GEOM_SPHERE:
        t.inertia[0] = t.inertia[1] = t.inertia[2] = 2 * mass * size[0] * size[0] / 5;
        break;
      case GEOM_CAPSULE: {
        double height = 2 * size[1], radius = size[0];
        double sphere_mass = mass * 4 * radius / (4 * radius + 3 * height);
        double cylinder_mass = mass - sphere_mass;
        t.inertia[0] = t.inertia[1] = cylinder_mass * (3 * radius * radius + height * height) / 12;
        t.inertia[2] = cylinder_mass * radius * radius / 2;
        double sphere_inertia = 2 * sphere_mass * radius * radius / 5;
        t.inertia[0] += sphere_inertia + sphere_mass * height * (3 * radius + 2 * height) / 8;
        t.inertia[1] += sphere_inertia + sphere_mass * height * (3 * radius + 2 * height) / 8;
        t.inertia[2] += sphere_inertia;
        break;
      }
      case GEOM_CYLINDER: {
        double height = 2 * size[1], radius = size[0];
        t.inertia[0] = t.inertia[1] = mass * (3 * radius * radius + height * height) / 12;
        t.inertia[2] = mass * radius * radius / 2;
        break;
      }
      case GEOM_BOX:
        t.inertia[0] = mass * (size[1] * size[1] + size[2] * size[2]) / 3;
        t.inertia[1] = mass * (size[0] * size[0] + size[2] * size[2]) / 3;
        t.inertia[2] = mass * (size[0] * size[0] + size[1] * size[1]) / 3;
        break;
      case GEOM_MESH:
        if (mesh) for (int i = 0; i < 3; i++) t.inertia[i] = mass * mesh->inertia_unit[i] / mesh->volume;
        break;
      default: break;
    }
    geom_tmp.push_back(t);
  }

  void parse_body(const XmlNode& node, int parent, std::string childclass) {
    int body = (int)m.body_parentid.size();
    bool world = (parent < 0);
    if (auto* cc = node.attr("childclass")) childclass = *cc;
    m.body_parentid.push_back(world ? 0 : parent);
    m.body_names.push_back(world ? "world" : (node.attr("name") ? *node.attr("name") : ""));
    double pos[3] = {0, 0, 0}, quat[4] = {1, 0, 0, 0};
    if (!world) {
      Getter g{&node, nullptr};
      auto p = g.reals("pos");
      if (p.size() == 3) for (int i = 0; i < 3; i++) pos[i] = p[i];
      orientation(g, quat);
    }
    push_n(m.body_pos, pos, 3);
    push_n(m.body_quat, quat, 4);
    m.body_gravcomp.push_back(world ? 0.0 : Getter{&node, nullptr}.real("gravcomp", 0.0));
    m.body_jntadr.push_back(-1);
    m.body_jntnum.push_back(0);
    m.body_dofadr.push_back(-1);
    m.body_dofnum.push_back(0);
    m.body_geomadr.push_back(-1);
    m.body_geomnum.push_back(0);
    inertial_tmp.emplace_back();
    // joints first (they must be contiguous per body), then geoms, then child bodies: MuJoCo
    // numbers joints/geoms depth-first in document order, which this reproduces because each
    // body's own elements are added before recursing.
    for (auto& ch : node.children) {
      if (ch->tag == "joint" || ch->tag == "freejoint") {
        if (world) throw std::runtime_error("mjcf: joint in worldbody");
        if (m.body_jntnum[body] == 0) {
          m.body_jntadr[body] = (int)m.jnt_type.size();
          m.body_dofadr[body] = (int)m.dof_bodyid.size();
        }
        int nv0 = (int)m.dof_bodyid.size();
        add_joint(*ch, body, childclass, ch->tag == "freejoint");
        m.body_jntnum[body]++;
        m.body_dofnum[body] += (int)m.dof_bodyid.size() - nv0;
      }
    }
    for (auto& ch : node.children) {
      if (ch->tag == "geom") {
        if (m.body_geomnum[body] == 0) m.body_geomadr[body] = (int)m.geom_type.size();
        add_geom(*ch, body, childclass);
        m.body_geomnum[body]++;
      } else if (ch->tag == "site") {
        Getter g{ch.get(), nullptr};
        m.site_names.push_back(g.str("name", ""));
        m.site_bodyid.push_back(body);
        double sp[3] = {0, 0, 0};
        auto p = g.reals("pos");
        if (p.size() == 3) for (int i = 0; i < 3; i++) sp[i] = p[i];
        push_n(m.site_pos, sp, 3);
      } else if (ch->tag == "inertial") {
        Getter g{ch.get(), nullptr};
        InertialTmp& it = inertial_tmp[body];
        it.given = true;
        auto p = g.reals("pos");
        for (int i = 0; i < 3; i++) it.pos[i] = p.size() == 3 ? p[i] : 0;
        orientation(g, it.quat);
        it.mass = g.real("mass", 0);
        if (g.has("diaginertia")) {
          auto d = g.reals("diaginertia");
          for (int i = 0; i < 3; i++) it.inertia[i] = d[i];
        } else if (g.has("fullinertia")) {
          auto f = g.reals("fullinertia");  // xx yy zz xy xz yz
          double A[9] = {f[0], f[3], f[4], f[3], f[1], f[5], f[4], f[5], f[2]};
          double ev[3], R[9];
          hm::eig3(A, ev, R);
          for (int i = 0; i < 3; i++) it.inertia[i] = ev[i];
          hm::mat2quat(it.quat, R);
        } else {
          throw std::runtime_error("mjcf: <inertial> needs diaginertia or fullinertia");
        }
      }
    }
    for (auto& ch : node.children)
      if (ch->tag == "body") parse_body(*ch, body, childclass);
  }

  // ---- inertia from geoms (MuJoCo: mjCBody::GeomFrame / inertiafromgeom="auto") --------------
  void finish_inertia() {
    int nbody = (int)m.body_parentid.size();
    m.body_mass.assign(nbody, 0);
    m.body_inertia.assign(3 * nbody, 0);
    m.body_ipos.assign(3 * nbody, 0);
    m.body_iquat.assign(4 * nbody, 0);
    for (int b = 0; b < nbody; b++) {
      double* ipos = &m.body_ipos[3 * b];
      double* iquat = &m.body_iquat[4 * b];
      iquat[0] = 1;
      if (b == 0) continue;
      if (inertial_tmp[b].given) {
        const InertialTmp& it = inertial_tmp[b];
        m.body_mass[b] = it.mass;
        for (int i = 0; i < 3; i++) { ipos[i] = it.pos[i]; m.body_inertia[3 * b + i] = it.inertia[i]; }
        for (int i = 0; i < 4; i++) iquat[i] = it.quat[i];
        continue;
      }
      int g0 = m.body_geomadr[b], ng = m.body_geomnum[b];
      std::vector<int> gs;
      for (int g = g0; g < g0 + ng; g++)
        if (geom_tmp[g].mass > 0) gs.push_back(g);
      if (gs.empty()) continue;
      if (gs.size() == 1) {
        int g = gs[0];
        m.body_mass[b] = geom_tmp[g].mass;
        for (int i = 0; i < 3; i++) { ipos[i] = m.geom_pos[3 * g + i]; m.body_inertia[3 * b + i] = geom_tmp[g].inertia[i]; }
        for (int i = 0; i < 4; i++) iquat[i] = m.geom_quat[4 * g + i];
        continue;
      }
      double mass = 0, com[3] = {0, 0, 0};
      for (int g : gs) {
        mass += geom_tmp[g].mass;
        for (int i = 0; i < 3; i++) com[i] += geom_tmp[g].mass * m.geom_pos[3 * g + i];
      }
      for (int i = 0; i < 3; i++) com[i] /= mass;
      double I[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int g : gs) {
        double R[9];
        hm::quat2mat(R, &m.geom_quat[4 * g]);
        const double* gi = geom_tmp[g].inertia;
        double d[3];
        for (int i = 0; i < 3; i++) d[i] = m.geom_pos[3 * g + i] - com[i];
        double gm = geom_tmp[g].mass, d2 = hm::dot3(d, d);
        for (int r = 0; r < 3; r++)
          for (int c = 0; c < 3; c++) {
            double v = 0;
            for (int k = 0; k < 3; k++) v += R[3 * r + k] * gi[k] * R[3 * c + k];
            v += gm * ((r == c ? d2 : 0.0) - d[r] * d[c]);
            I[3 * r + c] += v;
          }
      }
      double ev[3], R[9];
      hm::eig3(I, ev, R);
      m.body_mass[b] = mass;
      for (int i = 0; i < 3; i++) { ipos[i] = com[i]; m.body_inertia[3 * b + i] = ev[i]; }
      hm::mat2quat(iquat, R);
    }
  }

  // ---- actuators ----------------------------------------------------------------------------
  void parse_actuators(const XmlNode& node) {
    for (auto& ch : node.children) {
      const std::string& tag = ch->tag;
      if (tag != "motor" && tag != "position" && tag != "general")
        throw std::runtime_error("mjcf: actuator <" + tag + "> is outside the supported subset");
      Getter g = getter(*ch, tag, "");
      std::string jn = g.str("joint", "");
      int jid = -1;
      for (size_t j = 0; j < m.jnt_names.size(); j++)
        if (m.jnt_names[j] == jn) jid = (int)j;
      if (jid < 0) throw std::runtime_error("mjcf: actuator joint '" + jn + "' not found");
      if (m.jnt_type[jid] == JNT_FREE) throw std::runtime_error("mjcf: actuator on a free joint unsupported");
      m.actuator_names.push_back(g.str("name", ""));
      m.actuator_trnid.push_back(jid);
      m.actuator_dofadr.push_back(m.jnt_dofadr[jid]);
      auto gear = g.reals("gear");
      m.actuator_gear.push_back(gear.empty() ? 1.0 : gear[0]);
      double cr[2] = {0, 0}, fr[2] = {0, 0};
      auto c = g.reals("ctrlrange");
      bool has_cr = c.size() == 2;
      if (has_cr) { cr[0] = c[0]; cr[1] = c[1]; }
      auto f = g.reals("forcerange");
      bool has_fr = f.size() == 2;
      if (has_fr) { fr[0] = f[0]; fr[1] = f[1]; }
      std::string cl = g.str("ctrllimited", "auto"), fl = g.str("forcelimited", "auto");
      m.actuator_ctrllimited.push_back(cl == "auto" ? (autolimits && has_cr) : parse_bool(cl));
      m.actuator_forcelimited.push_back(fl == "auto" ? (autolimits && has_fr) : parse_bool(fl));
      push_n(m.actuator_ctrlrange, cr, 2);
      push_n(m.actuator_forcerange, fr, 2);
      double gain = 1.0, bias[3] = {0, 0, 0};
      int biastype = 0;  // 0 none, 1 affine
      if (tag == "position") {
        double kp = g.real("kp", 1.0), kv = g.real("kv", 0.0);
        gain = kp; bias[1] = -kp; bias[2] = -kv; biastype = 1;
      } else if (tag == "general") {
        auto gp = g.reals("gainprm");
        if (!gp.empty()) gain = gp[0];
        auto bp = g.reals("biasprm");
        for (size_t i = 0; i < bp.size() && i < 3; i++) bias[i] = bp[i];
        biastype = g.str("biastype", "none") == "affine" ? 1 : 0;
      }
      m.actuator_gainprm.push_back(gain);
      push_n(m.actuator_biasprm, bias, 3);
      m.actuator_biastype.push_back(biastype);
    }
  }

  // ---- tree bookkeeping ---------------------------------------------------------------------
  void finish_tree() {
    int nbody = (int)m.body_parentid.size(), nv = (int)m.dof_bodyid.size();
    if (nbody > 64) throw std::runtime_error("model: more than 64 bodies unsupported");
    if (nv > 64) throw std::runtime_error("model: more than 64 dofs unsupported");
    m.body_rootid.assign(nbody, 0);
    m.body_weldid.assign(nbody, 0);
    m.body_depth.assign(nbody, 0);
    int maxdepth = 0;
    for (int b = 1; b < nbody; b++) {
      int p = m.body_parentid[b];
      m.body_rootid[b] = (p == 0) ? b : m.body_rootid[p];
      m.body_weldid[b] = m.body_jntnum[b] > 0 ? b : m.body_weldid[p];
      m.body_depth[b] = m.body_depth[p] + 1;
      maxdepth = std::max(maxdepth, m.body_depth[b]);
    }
    std::vector<uint64_t> sub(nbody, 0);
    for (int b = 0; b < nbody; b++) sub[b] = 1ull << b;
    for (int b = nbody - 1; b >= 1; b--) sub[m.body_parentid[b]] |= sub[b];
    m.body_subtreemask_lo.resize(nbody);
    m.body_subtreemask_hi.resize(nbody);
    for (int b = 0; b < nbody; b++) {
      m.body_subtreemask_lo[b] = (int32_t)(uint32_t)(sub[b] & 0xffffffffull);
      m.body_subtreemask_hi[b] = (int32_t)(uint32_t)(sub[b] >> 32);
    }
    // dof parents
    m.dof_parentid.assign(nv, -1);
    for (int d = 0; d < nv; d++) {
      int b = m.dof_bodyid[d];
      if (d > m.body_dofadr[b]) { m.dof_parentid[d] = d - 1; continue; }
      int p = m.body_parentid[b];
      while (p > 0 && m.body_dofnum[p] == 0) p = m.body_parentid[p];
      if (p > 0) m.dof_parentid[d] = m.body_dofadr[p] + m.body_dofnum[p] - 1;
    }
    m.dof_ancmask_lo.resize(nv);
    m.dof_ancmask_hi.resize(nv);
    for (int d = 0; d < nv; d++) {
      uint64_t mask = 0;
      for (int j = d; j >= 0; j = m.dof_parentid[j]) mask |= 1ull << j;
      m.dof_ancmask_lo[d] = (int32_t)(uint32_t)(mask & 0xffffffffull);
      m.dof_ancmask_hi[d] = (int32_t)(uint32_t)(mask >> 32);
    }
    // last dof of each dof's kinematic tree (bodies are in depth-first order, so a tree's dofs are one index range)
    m.dof_treelast.assign(nv, 0);
    {
      std::vector<int> root(nv), last(nv, -1);
      for (int d = 0; d < nv; d++) {
        int r = d;
        while (m.dof_parentid[r] >= 0) r = m.dof_parentid[r];
        root[d] = r;
        if (d > last[r]) last[r] = d;
      }
      for (int d = 0; d < nv; d++) m.dof_treelast[d] = last[root[d]];
    }
    treedepth_ = maxdepth;
  }
  int treedepth_ = 0;

  // ---- static collision pair list (MuJoCo broadphase filters that do not depend on state) -----
  static int pair_maxcontacts(int t1, int t2) {
    if (t1 == GEOM_PLANE) {
      switch (t2) {
        case GEOM_SPHERE: return 1;
        case GEOM_CAPSULE: return 2;
        case GEOM_CYLINDER: return 4;
        case GEOM_BOX: return 4;
        case GEOM_MESH: return 1;  // deepest vertex only (no hull graph; see step_convex.cuh)
        default: return 4;
      }
    }
    if (t1 == GEOM_SPHERE || t2 == GEOM_SPHERE) return 1;
    if (t1 == GEOM_CAPSULE && t2 == GEOM_CAPSULE) return 1;
    if (t1 == GEOM_CAPSULE && t2 == GEOM_BOX) return 2;
    if (t1 == GEOM_BOX && t2 == GEOM_BOX) return 8;
    return 1;
  }

  void finish_pairs() {
    int nbody = (int)m.body_parentid.size();
    int maxcon = 0, maxefc = 0;
    for (int b1 = 0; b1 < nbody; b1++)
      for (int b2 = b1 + 1; b2 < nbody; b2++) {
        int w1 = m.body_weldid[b1], w2 = m.body_weldid[b2];
        if (w1 == w2) continue;  // same (welded) body; also covers two static bodies
        // parent-child filter, world excluded
        int pw1 = m.body_weldid[m.body_parentid[w1]], pw2 = m.body_weldid[m.body_parentid[w2]];
        if (w1 != 0 && w2 != 0 && (pw1 == w2 || pw2 == w1)) continue;
        for (int g1 = m.body_geomadr[b1]; g1 >= 0 && g1 < m.body_geomadr[b1] + m.body_geomnum[b1]; g1++)
          for (int g2 = m.body_geomadr[b2]; g2 >= 0 && g2 < m.body_geomadr[b2] + m.body_geomnum[b2]; g2++) {
            bool ok = (m.geom_contype[g1] & m.geom_conaffinity[g2]) || (m.geom_contype[g2] & m.geom_conaffinity[g1]);
            if (!ok) continue;
            int a = g1, b = g2;
            if (m.geom_type[a] > m.geom_type[b]) std::swap(a, b);
            if (m.geom_type[a] == GEOM_PLANE && m.geom_type[b] == GEOM_PLANE) continue;
            m.pair_geom1.push_back(a);
            m.pair_geom2.push_back(b);
            m.pair_slotadr.push_back(maxcon);
            int nc = pair_maxcontacts(m.geom_type[a], m.geom_type[b]);
            int condim = std::max(m.geom_condim[a], m.geom_condim[b]);
            if (m.geom_priority[a] != m.geom_priority[b])
              condim = m.geom_priority[a] > m.geom_priority[b] ? m.geom_condim[a] : m.geom_condim[b];
            // fail at load time, not silently inside a step: the narrow phase covers plane-{sphere, capsule, box,
            // cylinder, mesh} and the convex collider's pairs (mesh-anything convex, cylinder-cylinder, box-cylinder);
            // the constraint rows cover condim 1 and 3 (pyramidal)
            {
              int ta = m.geom_type[a], tb = m.geom_type[b];
              auto convex = [](int t) { return t == GEOM_SPHERE || t == GEOM_CAPSULE || t == GEOM_CYLINDER || t == GEOM_BOX || t == GEOM_MESH; };
              bool have = (ta == GEOM_PLANE && convex(tb)) ||
                          (convex(ta) && convex(tb) &&
                           (ta == GEOM_MESH || tb == GEOM_MESH || (ta == GEOM_CYLINDER && tb == GEOM_CYLINDER) ||
                            (ta == GEOM_CYLINDER && tb == GEOM_BOX) || (ta == GEOM_BOX && tb == GEOM_CYLINDER)));
              if (!have)
                throw std::runtime_error("unsupported collision pair: geoms '" + m.geom_names[a] + "' (type " + std::to_string(ta) +
                                         ") and '" + m.geom_names[b] + "' (type " + std::to_string(tb) +
                                         ") have no narrow-phase function in this build (disable the pair with contype/conaffinity)");
              if (condim != 1 && condim != 3)
                throw std::runtime_error("unsupported contact dimension " + std::to_string(condim) + " for geoms '" + m.geom_names[a] +
                                         "' / '" + m.geom_names[b] + "' (condim 1 and 3 are implemented)");
            }
            maxcon += nc;
            maxefc += nc * (condim == 1 ? 1 : 2 * (condim - 1));
          }
      }
    for (size_t j = 0; j < m.jnt_limited.size(); j++)
      if (m.jnt_limited[j]) maxefc += 1;
    maxcontact_ = maxcon;
    maxefc_ = maxefc;
  }
  int maxcontact_ = 0, maxefc_ = 0;

  // ---- qpos0 constants (MuJoCo: mj_setConst / set0) -----------------------------------------
  void finish_const() {
    int nbody = m.nbody(), nv = m.nv();
    m.dof_invweight0.assign(nv, 0);
    m.body_invweight0.assign(2 * nbody, 0);
    if (nv == 0) { m.opt_real[7] = 1; return; }
    // kinematics at qpos0
    std::vector<double> xpos(3 * nbody, 0), xquat(4 * nbody, 0), xmat(9 * nbody, 0), xipos(3 * nbody, 0), ximat(9 * nbody, 0);
    std::vector<double> xanchor(3 * m.njnt(), 0), xaxis(3 * m.njnt(), 0);
    xquat[0] = 1;
    hm::quat2mat(&xmat[0], &xquat[0]);
    hm::quat2mat(&ximat[0], &xquat[0]);
    for (int b = 1; b < nbody; b++) {
      int p = m.body_parentid[b];
      double pos[3], quat[4];
      hm::rot_vec_quat(pos, &m.body_pos[3 * b], &xquat[4 * p]);
      for (int i = 0; i < 3; i++) pos[i] += xpos[3 * p + i];
      hm::quat_mul(quat, &xquat[4 * p], &m.body_quat[4 * b]);
      for (int j = m.body_jntadr[b]; j >= 0 && j < m.body_jntadr[b] + m.body_jntnum[b]; j++) {
        // at qpos0 hinge/slide displacement is zero; free joint sets pose from qpos0
        if (m.jnt_type[j] == JNT_FREE) {
          for (int i = 0; i < 3; i++) pos[i] = m.qpos0[m.jnt_qposadr[j] + i];
          for (int i = 0; i < 4; i++) quat[i] = m.qpos0[m.jnt_qposadr[j] + 3 + i];
          hm::quat_normalize(quat);
        }
        double a[3];
        hm::rot_vec_quat(a, &m.jnt_pos[3 * j], quat);
        for (int i = 0; i < 3; i++) xanchor[3 * j + i] = a[i] + pos[i];
        hm::rot_vec_quat(&xaxis[3 * j], &m.jnt_axis[3 * j], quat);
      }
      hm::quat_normalize(quat);
      for (int i = 0; i < 3; i++) xpos[3 * b + i] = pos[i];
      for (int i = 0; i < 4; i++) xquat[4 * b + i] = quat[i];
      hm::quat2mat(&xmat[9 * b], quat);
      double ip[3], iq[4];
      hm::rot_vec_quat(ip, &m.body_ipos[3 * b], quat);
      for (int i = 0; i < 3; i++) xipos[3 * b + i] = pos[i] + ip[i];
      hm::quat_mul(iq, quat, &m.body_iquat[4 * b]);
      hm::quat2mat(&ximat[9 * b], iq);
    }
    // dense mass matrix by the Jacobian method: M = sum_b Jp^T m Jp + Jr^T I Jr (+ armature)
    std::vector<double> M(nv * nv, 0);
    std::vector<std::vector<double>> Jp(nbody, std::vector<double>(3 * nv, 0)), Jr(nbody, std::vector<double>(3 * nv, 0));
    for (int b = 1; b < nbody; b++) {
      // walk up the tree collecting dofs
      for (int a = b; a > 0; a = m.body_parentid[a]) {
        for (int j = m.body_jntadr[a]; j >= 0 && j < m.body_jntadr[a] + m.body_jntnum[a]; j++) {
          int d0 = m.jnt_dofadr[j];
          if (m.jnt_type[j] == JNT_SLIDE) {
            for (int i = 0; i < 3; i++) Jp[b][i * nv + d0] = xaxis[3 * j + i];
          } else if (m.jnt_type[j] == JNT_HINGE) {
            double r[3], c[3];
            for (int i = 0; i < 3; i++) r[i] = xipos[3 * b + i] - xanchor[3 * j + i];
            hm::cross(c, &xaxis[3 * j], r);
            for (int i = 0; i < 3; i++) { Jp[b][i * nv + d0] = c[i]; Jr[b][i * nv + d0] = xaxis[3 * j + i]; }
          } else if (m.jnt_type[j] == JNT_FREE) {
            // translational dofs: world axes; rotational dofs: body-frame axes of body a
            for (int i = 0; i < 3; i++) Jp[b][i * nv + d0 + i] = 1;
            for (int k = 0; k < 3; k++) {
              double ax[3] = {xmat[9 * a + k], xmat[9 * a + 3 + k], xmat[9 * a + 6 + k]};
              double r[3], c[3];
              for (int i = 0; i < 3; i++) r[i] = xipos[3 * b + i] - xpos[3 * a + i];
              hm::cross(c, ax, r);
              for (int i = 0; i < 3; i++) { Jp[b][i * nv + d0 + 3 + k] = c[i]; Jr[b][i * nv + d0 + 3 + k] = ax[i]; }
            }
          }
        }
      }
      // world-frame inertia
      double Iw[9];
      for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) {
          double v = 0;
          for (int k = 0; k < 3; k++) v += ximat[9 * b + 3 * r + k] * m.body_inertia[3 * b + k] * ximat[9 * b + 3 * c + k];
          Iw[3 * r + c] = v;
        }
      for (int i = 0; i < nv; i++)
        for (int j = 0; j < nv; j++) {
          double v = 0;
          for (int k = 0; k < 3; k++) v += m.body_mass[b] * Jp[b][k * nv + i] * Jp[b][k * nv + j];
          for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) v += Jr[b][r * nv + i] * Iw[3 * r + c] * Jr[b][c * nv + j];
          M[i * nv + j] += v;
        }
    }
    for (int i = 0; i < nv; i++) M[i * nv + i] += m.dof_armature[i];
    double mean = 0;
    for (int i = 0; i < nv; i++) mean += M[i * nv + i];
    m.opt_real[7] = mean / nv;
    // inverse via Cholesky
    std::vector<double> L(M), Minv(nv * nv, 0);
    for (int j = 0; j < nv; j++) {
      double s = L[j * nv + j];
      for (int k = 0; k < j; k++) s -= L[j * nv + k] * L[j * nv + k];
      if (s <= 0) throw std::runtime_error("model: mass matrix at qpos0 is not positive definite");
      L[j * nv + j] = std::sqrt(s);
      for (int i = j + 1; i < nv; i++) {
        double t = L[i * nv + j];
        for (int k = 0; k < j; k++) t -= L[i * nv + k] * L[j * nv + k];
        L[i * nv + j] = t / L[j * nv + j];
      }
    }
    for (int c = 0; c < nv; c++) {
      std::vector<double> x(nv, 0);
      x[c] = 1;
      for (int i = 0; i < nv; i++) {
        double t = x[i];
        for (int k = 0; k < i; k++) t -= L[i * nv + k] * x[k];
        x[i] = t / L[i * nv + i];
      }
      for (int i = nv - 1; i >= 0; i--) {
        double t = x[i];
        for (int k = i + 1; k < nv; k++) t -= L[k * nv + i] * x[k];
        x[i] = t / L[i * nv + i];
      }
      for (int i = 0; i < nv; i++) Minv[i * nv + c] = x[i];
    }
    // dof_invweight0 (free joints: average of the 3 translational / 3 rotational)
    for (int j = 0; j < m.njnt(); j++) {
      int d0 = m.jnt_dofadr[j];
      if (m.jnt_type[j] == JNT_FREE) {
        double t = (Minv[d0 * nv + d0] + Minv[(d0 + 1) * nv + d0 + 1] + Minv[(d0 + 2) * nv + d0 + 2]) / 3;
        double r = (Minv[(d0 + 3) * nv + d0 + 3] + Minv[(d0 + 4) * nv + d0 + 4] + Minv[(d0 + 5) * nv + d0 + 5]) / 3;
        for (int k = 0; k < 3; k++) { m.dof_invweight0[d0 + k] = t; m.dof_invweight0[d0 + 3 + k] = r; }
      } else {
        m.dof_invweight0[d0] = Minv[d0 * nv + d0];
      }
    }
    // body_invweight0 = trace(J Minv J^T)/3 for the translational (at body com) and rotational blocks
    for (int b = 1; b < nbody; b++) {
      if (m.body_weldid[b] == 0) continue;
      double tr[2] = {0, 0};
      for (int which = 0; which < 2; which++) {
        const std::vector<double>& J = which == 0 ? Jp[b] : Jr[b];
        for (int k = 0; k < 3; k++) {
          for (int i = 0; i < nv; i++) {
            if (J[k * nv + i] == 0) continue;
            for (int j = 0; j < nv; j++) tr[which] += J[k * nv + i] * Minv[i * nv + j] * J[k * nv + j];
          }
        }
      }
      m.body_invweight0[2 * b] = std::max(kMinVal, tr[0] / 3);
      m.body_invweight0[2 * b + 1] = std::max(kMinVal, tr[1] / 3);
    }
  }

  // ---- top level ------------------------------------------------------------------------------
  Model compile(const XmlNode& root, double timestep_override) {
    if (root.tag != "mujoco") throw std::runtime_error("mjcf: root element must be <mujoco>");
    for (auto& ch : root.children) {
      if (ch->tag == "compiler") {
        if (auto* a = ch->attr("angle")) angle_degree = (*a != "radian");
        if (auto* a = ch->attr("eulerseq")) eulerseq = *a;
        if (auto* a = ch->attr("autolimits")) autolimits = parse_bool(*a);
        if (auto* a = ch->attr("meshdir")) meshdir = *a;
        if (auto* a = ch->attr("coordinate"))
          if (*a != "local") throw std::runtime_error("mjcf: only coordinate=local is supported");
      } else if (ch->tag == "option") {
        if (auto* a = ch->attr("timestep")) opt_timestep = parse_reals(*a)[0];
        if (auto* a = ch->attr("gravity")) { auto v = parse_reals(*a); for (int i = 0; i < 3; i++) gravity[i] = v[i]; }
        if (auto* a = ch->attr("iterations")) opt_iterations = (int)parse_reals(*a)[0];
        if (auto* a = ch->attr("ls_iterations")) opt_ls_iterations = (int)parse_reals(*a)[0];
        if (auto* a = ch->attr("tolerance")) opt_tolerance = parse_reals(*a)[0];
        if (auto* a = ch->attr("ls_tolerance")) opt_ls_tolerance = parse_reals(*a)[0];
        if (auto* a = ch->attr("impratio")) opt_impratio = parse_reals(*a)[0];
        if (auto* a = ch->attr("integrator"))
          if (*a != "Euler") throw std::runtime_error("mjcf: only the Euler integrator is supported");
        if (auto* a = ch->attr("cone"))
          if (*a != "pyramidal") throw std::runtime_error("mjcf: only pyramidal cones are supported");
        if (auto* a = ch->attr("solver"))
          if (*a != "Newton") throw std::runtime_error("mjcf: only the Newton solver is supported");
      }
    }
    // dynamics the step does not model must not be dropped silently
    for (auto& ch : root.children)
      if ((ch->tag == "tendon" || ch->tag == "equality") && !ch->children.empty())
        throw std::runtime_error("mjcf: <" + ch->tag + "> elements are outside the supported subset");
    for (auto& ch : root.children)
      if (ch->tag == "default") parse_default(*ch, nullptr, true);
    for (auto& ch : root.children)
      if (ch->tag == "asset") parse_assets(*ch);
    const XmlNode* world = nullptr;
    for (auto& ch : root.children)
      if (ch->tag == "worldbody") { world = ch.get(); break; }  // compile_mjcf_string merged the others into it
    if (!world) throw std::runtime_error("mjcf: no <worldbody>");
    parse_body(*world, -1, "");
    for (auto& ch : root.children)
      if (ch->tag == "actuator") parse_actuators(*ch);
    finish_inertia();
    finish_tree();
    finish_pairs();
    int nq = (int)m.qpos0.size(), nv = (int)m.dof_bodyid.size();
    m.dims = {nq, nv, (int)m.actuator_trnid.size(), (int)m.body_parentid.size(), (int)m.jnt_type.size(),
              (int)m.geom_type.size(), (int)m.pair_geom1.size(), (int)(m.mesh_vert.size() / 3), maxcontact_, maxefc_,
              treedepth_};
    m.opt_int = {opt_iterations, opt_ls_iterations, 0};
    double ts = timestep_override > 0 ? timestep_override : opt_timestep;
    m.opt_real = {ts, gravity[0], gravity[1], gravity[2], opt_tolerance, opt_ls_tolerance, opt_impratio, 1.0};
    finish_const();
    return std::move(m);
  }
};

template <class T>
void put_field(std::vector<uint8_t>& out, const char* name, int32_t dtype, const std::vector<T>& v) {
  char nm[32];
  std::memset(nm, 0, sizeof(nm));
  std::strncpy(nm, name, 31);
  out.insert(out.end(), nm, nm + 32);
  const uint8_t* p = reinterpret_cast<const uint8_t*>(&dtype);
  out.insert(out.end(), p, p + 4);
  int64_t cnt = (int64_t)v.size();
  p = reinterpret_cast<const uint8_t*>(&cnt);
  out.insert(out.end(), p, p + 8);
  p = reinterpret_cast<const uint8_t*>(v.data());
  out.insert(out.end(), p, p + sizeof(T) * v.size());
}

}  // namespace

std::vector<uint8_t> Model::serialize() const {
  std::vector<uint8_t> out;
  const char magic[8] = {'T', 'O', 'K', 'P', 'M', 'D', 'L', '1'};
  out.insert(out.end(), magic, magic + 8);
#define X(name) put_field(out, #name, 0, name);
  TOKP_MODEL_INT_FIELDS(X)
#undef X
#define X(name) put_field(out, #name, 1, name);
  TOKP_MODEL_REAL_FIELDS(X)
#undef X
  return out;
}

Model compile_mjcf_string(const std::string& xml, const std::string& base_dir, double timestep) {
  XmlParser parser(xml);
  auto root = parser.parse();
  // several <worldbody> elements (resolved <include>s) are one world body, children in document order
  XmlNode* first = nullptr;
  for (auto& ch : root->children) {
    if (ch->tag != "worldbody") continue;
    if (!first) { first = ch.get(); continue; }
    for (auto& g : ch->children) first->children.push_back(std::move(g));
    ch->children.clear();
    ch->tag = "worldbody_merged";
  }
  Compiler c;
  c.base_dir = base_dir;
  return c.compile(*root, timestep);
}

Model compile_mjcf_file(const std::string& path, double timestep) {
  std::ifstream f(path, std::ios::binary);
  if (!f) throw std::runtime_error("mjcf: cannot open '" + path + "'");
  std::stringstream ss;
  ss << f.rdbuf();
  std::string dir = ".";
  size_t slash = path.find_last_of('/');
  if (slash != std::string::npos) dir = path.substr(0, slash);
  return compile_mjcf_string(ss.str(), dir, timestep);
}

}  // namespace tokp
