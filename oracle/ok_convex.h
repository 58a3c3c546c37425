// ORACLE (test infrastructure only).  Convex-convex and plane-convex collision.
//
// MuJoCo's general convex collider (engine_collision_convex.c: mjc_Convex, mjc_PlaneConvex, mjccd_support,
// mjccd_center) calls libccd's Minkowski Portal Refinement, ccdMPRPenetration (libccd src/mpr.c; algorithm:
// G. Snethen, "XenoCollide", Game Programming Gems 7, 2008), with mpr_tolerance = 1e-6 and at most 50 refinement
// iterations, on the two geoms inflated by margin/2 each; the contact is {dist = margin - depth, normal = dir,
// pos = midpoint of the witness points}.  Neither MuJoCo nor libccd is in /root/reference (external, un-vendored,
// version not pinned: SURVEY.md 8(c)), so this file restates the published algorithm; parity with MuJoCo's bits is
// UNPINNED for these pair types.  Call sites on the reference path: every mj_step / mj_stepSkip of the Franka
// scenes (src/Differentiator/Differentiator.cpp:69,118,...; src/Optimiser/iLQR.cpp:222,914).
//
// Mesh geoms collide with their convex hull; the support function runs over all (merged) vertices, which has the
// same maximiser as the hull.  Known deviations, flagged where they can matter:
//   * mjc_PlaneConvex adds up to three hull neighbours of the deepest vertex as extra contacts; no hull graph is
//     built here, so a mesh touching a plane yields its deepest vertex only;
//   * mjc_fixNormal (smooth geoms: sphere / capsule / ellipsoid / cylinder) is not applied.
#pragma once
#include "ok_sim.h"

namespace ok {

constexpr double kMprTolerance = 1e-6;
constexpr int kMprIterations = 50;
constexpr double kCcdEps = 2.220446049250313e-16;  // DBL_EPSILON (libccd CCD_EPS in double builds)

template <class R>
struct ConvexObj {
  const Model* m;
  int geom, type;
  const R* pos;   // geom_xpos
  const R* mat;   // geom_xmat (row-major, columns = geom axes in the world)
  R margin;       // half of the pair margin
};

template <class R> inline bool ccd_zero(R x) { using std::fabs; return fabs(x) < R(kCcdEps); }

// mjccd_support: farthest point of the (inflated) geom in the world direction `dir` (unit length)
template <class R>
void convex_support(const ConvexObj<R>& o, const R* dir, R* res) {
  using std::sqrt;
  const double* size = &o.m->geom_size[3 * o.geom];
  R ld[3], lp[3] = {R(0.0), R(0.0), R(0.0)};
  for (int k = 0; k < 3; k++) ld[k] = o.mat[k] * dir[0] + o.mat[3 + k] * dir[1] + o.mat[6 + k] * dir[2];
  switch (o.type) {
    case GEOM_SPHERE:
      for (int k = 0; k < 3; k++) lp[k] = ld[k] * R(size[0]);
      break;
    case GEOM_CAPSULE:
      for (int k = 0; k < 3; k++) lp[k] = ld[k] * R(size[0]);
      lp[2] = lp[2] + (ld[2] >= R(0.0) ? R(size[1]) : R(-size[1]));
      break;
    case GEOM_CYLINDER: {
      R r2 = ld[0] * ld[0] + ld[1] * ld[1];
      if (r2 > R(kMinVal) * R(kMinVal)) {
        R s = R(size[0]) / sqrt(r2);
        lp[0] = ld[0] * s; lp[1] = ld[1] * s;
      }
      lp[2] = ld[2] >= R(0.0) ? R(size[1]) : R(-size[1]);
      break;
    }
    case GEOM_BOX:
      for (int k = 0; k < 3; k++) lp[k] = ld[k] >= R(0.0) ? R(size[k]) : R(-size[k]);
      break;
    case GEOM_MESH: {
      const double* v = &o.m->mesh_vert[3 * o.m->geom_meshadr[o.geom]];
      int n = o.m->geom_meshnum[o.geom], best = 0;
      R bd = R(v[0]) * ld[0] + R(v[1]) * ld[1] + R(v[2]) * ld[2];  // NOLINT
      for (int i = 1; i < n; i++) {
        R dd = R(v[3 * i]) * ld[0] + R(v[3 * i + 1]) * ld[1] + R(v[3 * i + 2]) * ld[2];
        if (dd > bd) { bd = dd; best = i; }
      }
      for (int k = 0; k < 3; k++) lp[k] = R(v[3 * best + k]);
      break;
    }
    default: break;
  }
  for (int k = 0; k < 3; k++)
    res[k] = o.pos[k] + o.mat[3 * k] * lp[0] + o.mat[3 * k + 1] * lp[1] + o.mat[3 * k + 2] * lp[2] + dir[k] * o.margin;
}

// a point of the Minkowski difference with its witnesses on both geoms
template <class R>
struct MprPoint { R v[3], v1[3], v2[3]; };

template <class R>
void mpr_support(const ConvexObj<R>& a, const ConvexObj<R>& b, const R* dir, MprPoint<R>& p) {
  R nd[3] = {-dir[0], -dir[1], -dir[2]};
  convex_support(a, dir, p.v1);
  convex_support(b, nd, p.v2);
  for (int k = 0; k < 3; k++) p.v[k] = p.v1[k] - p.v2[k];
}

// closest point of triangle (a, b, c) to the origin (C. Ericson, Real-Time Collision Detection, 5.1.5)
template <class R>
void closest_on_triangle_to_origin(const R* a, const R* b, const R* c, R* out) {
  R ab[3], ac[3], ap[3], bp[3], cp[3];
  for (int k = 0; k < 3; k++) { ab[k] = b[k] - a[k]; ac[k] = c[k] - a[k]; ap[k] = -a[k]; bp[k] = -b[k]; cp[k] = -c[k]; }
  R d1 = dot3(ab, ap), d2 = dot3(ac, ap);
  if (d1 <= R(0.0) && d2 <= R(0.0)) { copy(out, a, 3); return; }
  R d3 = dot3(ab, bp), d4 = dot3(ac, bp);
  if (d3 >= R(0.0) && d4 <= d3) { copy(out, b, 3); return; }
  R vc = d1 * d4 - d3 * d2;
  if (vc <= R(0.0) && d1 >= R(0.0) && d3 <= R(0.0)) {
    R v = d1 / (d1 - d3);
    for (int k = 0; k < 3; k++) out[k] = a[k] + v * ab[k];
    return;
  }
  R d5 = dot3(ab, cp), d6 = dot3(ac, cp);
  if (d6 >= R(0.0) && d5 <= d6) { copy(out, c, 3); return; }
  R vb = d5 * d2 - d1 * d6;
  if (vb <= R(0.0) && d2 >= R(0.0) && d6 <= R(0.0)) {
    R w = d2 / (d2 - d6);
    for (int k = 0; k < 3; k++) out[k] = a[k] + w * ac[k];
    return;
  }
  R va = d3 * d6 - d5 * d4;
  if (va <= R(0.0) && (d4 - d3) >= R(0.0) && (d5 - d6) >= R(0.0)) {
    R w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    for (int k = 0; k < 3; k++) out[k] = b[k] + w * (c[k] - b[k]);
    return;
  }
  R denom = R(1.0) / (va + vb + vc);
  R v = vb * denom, w = vc * denom;
  for (int k = 0; k < 3; k++) out[k] = a[k] + ab[k] * v + ac[k] * w;
}

template <class R>
struct MprPortal { MprPoint<R> p[4]; };  // p[0] = interior point v0, p[1..3] = portal triangle

template <class R>
inline void mpr_portal_dir(const MprPortal<R>& s, R* dir) {
  R e1[3], e2[3];
  for (int k = 0; k < 3; k++) { e1[k] = s.p[2].v[k] - s.p[1].v[k]; e2[k] = s.p[3].v[k] - s.p[1].v[k]; }
  cross(dir, e1, e2);
  normalize3(dir);
}

template <class R>
inline bool mpr_reach_tolerance(const MprPortal<R>& s, const MprPoint<R>& v4, const R* dir) {
  R dv4 = dot3(v4.v, dir);
  R d1 = dv4 - dot3(s.p[1].v, dir), d2 = dv4 - dot3(s.p[2].v, dir), d3 = dv4 - dot3(s.p[3].v, dir);
  R mn = d1 < d2 ? d1 : d2;
  if (d3 < mn) mn = d3;
  return mn <= R(kMprTolerance);
}

template <class R>
inline void mpr_expand(MprPortal<R>& s, const MprPoint<R>& v4) {
  R v4v0[3];
  cross(v4v0, v4.v, s.p[0].v);
  if (dot3(s.p[1].v, v4v0) > R(0.0)) {
    if (dot3(s.p[2].v, v4v0) > R(0.0)) s.p[1] = v4; else s.p[3] = v4;
  } else {
    if (dot3(s.p[3].v, v4v0) > R(0.0)) s.p[2] = v4; else s.p[1] = v4;
  }
}

// ccdMPRPenetration: returns 1 and fills depth / dir / pos when the inflated geoms intersect, else 0
template <class R>
int mpr_penetration(const ConvexObj<R>& a, const ConvexObj<R>& b, R* depth, R* dir_out, R* pos_out) {
  using std::sqrt;
  MprPortal<R> s;
  R dir[3], va[3], vb[3];
  // ---- portal discovery
  for (int k = 0; k < 3; k++) { s.p[0].v1[k] = a.pos[k]; s.p[0].v2[k] = b.pos[k]; s.p[0].v[k] = a.pos[k] - b.pos[k]; }
  if (ccd_zero(s.p[0].v[0]) && ccd_zero(s.p[0].v[1]) && ccd_zero(s.p[0].v[2])) s.p[0].v[0] = s.p[0].v[0] + R(10.0 * kCcdEps);
  for (int k = 0; k < 3; k++) dir[k] = -s.p[0].v[k];
  normalize3(dir);
  mpr_support(a, b, dir, s.p[1]);
  R d = dot3(s.p[1].v, dir);
  if (ccd_zero(d) || d < R(0.0)) return 0;
  cross(dir, s.p[0].v, s.p[1].v);
  if (ccd_zero(dir[0]) && ccd_zero(dir[1]) && ccd_zero(dir[2])) {
    // the origin lies on the ray v0 -> v1
    if (ccd_zero(s.p[1].v[0]) && ccd_zero(s.p[1].v[1]) && ccd_zero(s.p[1].v[2])) {  // touching contact
      *depth = R(0.0);
      for (int k = 0; k < 3; k++) { dir_out[k] = R(0.0); pos_out[k] = R(0.5) * (s.p[1].v1[k] + s.p[1].v2[k]); }
      return 1;
    }
    for (int k = 0; k < 3; k++) { dir_out[k] = s.p[1].v[k]; pos_out[k] = R(0.5) * (s.p[1].v1[k] + s.p[1].v2[k]); }
    { R nn = sqrt(dot3(dir_out, dir_out)); *depth = nn; for (int k = 0; k < 3; k++) dir_out[k] = dir_out[k] / nn; }
    return 1;
  }
  normalize3(dir);
  mpr_support(a, b, dir, s.p[2]);
  d = dot3(s.p[2].v, dir);
  if (ccd_zero(d) || d < R(0.0)) return 0;
  for (int k = 0; k < 3; k++) { va[k] = s.p[1].v[k] - s.p[0].v[k]; vb[k] = s.p[2].v[k] - s.p[0].v[k]; }
  cross(dir, va, vb);
  normalize3(dir);
  if (dot3(dir, s.p[0].v) > R(0.0)) {  // orient the triangle away from v0
    MprPoint<R> t = s.p[1]; s.p[1] = s.p[2]; s.p[2] = t;
    for (int k = 0; k < 3; k++) dir[k] = -dir[k];
  }
  for (int guard = 0;; guard++) {
    if (guard > 100) return 0;
    mpr_support(a, b, dir, s.p[3]);
    d = dot3(s.p[3].v, dir);
    if (ccd_zero(d) || d < R(0.0)) return 0;
    bool cont = false;
    cross(va, s.p[1].v, s.p[3].v);
    d = dot3(va, s.p[0].v);
    if (d < R(0.0) && !ccd_zero(d)) { s.p[2] = s.p[3]; cont = true; }
    if (!cont) {
      cross(va, s.p[3].v, s.p[2].v);
      d = dot3(va, s.p[0].v);
      if (d < R(0.0) && !ccd_zero(d)) { s.p[1] = s.p[3]; cont = true; }
    }
    if (!cont) break;
    for (int k = 0; k < 3; k++) { va[k] = s.p[1].v[k] - s.p[0].v[k]; vb[k] = s.p[2].v[k] - s.p[0].v[k]; }
    cross(dir, va, vb);
    normalize3(dir);
  }
  // ---- portal refinement until the portal lies beyond the origin
  MprPoint<R> v4;
  for (int guard = 0;; guard++) {
    mpr_portal_dir(s, dir);
    d = dot3(dir, s.p[1].v);
    if (ccd_zero(d) || d > R(0.0)) break;  // portal encapsulates the origin
    mpr_support(a, b, dir, v4);
    d = dot3(v4.v, dir);
    if (!(ccd_zero(d) || d > R(0.0)) || mpr_reach_tolerance(s, v4, dir) || guard > 1000) return 0;
    mpr_expand(s, v4);
  }
  // ---- penetration: push the portal to the surface of the Minkowski difference
  for (int it = 0;; it++) {
    mpr_portal_dir(s, dir);
    mpr_support(a, b, dir, v4);
    if (mpr_reach_tolerance(s, v4, dir) || it > kMprIterations) {
      R w[3];
      closest_on_triangle_to_origin(s.p[1].v, s.p[2].v, s.p[3].v, w);
      *depth = sqrt(dot3(w, w));
      if (ccd_zero(w[0]) && ccd_zero(w[1]) && ccd_zero(w[2])) copy(w, dir, 3);
      normalize3(w);
      copy(dir_out, w, 3);
      // witness position: barycentric weights of the origin ray in the final portal
      R b4[4], t[3];
      cross(t, s.p[1].v, s.p[2].v); b4[0] = dot3(t, s.p[3].v);
      cross(t, s.p[3].v, s.p[2].v); b4[1] = dot3(t, s.p[0].v);
      cross(t, s.p[0].v, s.p[1].v); b4[2] = dot3(t, s.p[3].v);
      cross(t, s.p[2].v, s.p[1].v); b4[3] = dot3(t, s.p[0].v);
      R sum = b4[0] + b4[1] + b4[2] + b4[3];
      if (ccd_zero(sum) || sum < R(0.0)) {
        b4[0] = R(0.0);
        cross(t, s.p[2].v, s.p[3].v); b4[1] = dot3(t, dir);
        cross(t, s.p[3].v, s.p[1].v); b4[2] = dot3(t, dir);
        cross(t, s.p[1].v, s.p[2].v); b4[3] = dot3(t, dir);
        sum = b4[1] + b4[2] + b4[3];
      }
      R inv = R(1.0) / sum;
      for (int k = 0; k < 3; k++) {
        R p1 = R(0.0), p2 = R(0.0);
        for (int i = 0; i < 4; i++) { p1 = p1 + b4[i] * s.p[i].v1[k]; p2 = p2 + b4[i] * s.p[i].v2[k]; }
        pos_out[k] = R(0.5) * (p1 + p2) * inv;
      }
      return 1;
    }
    mpr_expand(s, v4);
  }
}

// mjc_Convex (single contact; mjENBL_MULTICCD is off by default)
template <class R>
int convex_convex(const Model* m, Contact<R>* con, R margin, int g1, const R* pos1, const R* mat1, int g2, const R* pos2,
                  const R* mat2) {
  ConvexObj<R> a{m, g1, m->geom_type[g1], pos1, mat1, R(0.5) * margin};
  ConvexObj<R> b{m, g2, m->geom_type[g2], pos2, mat2, R(0.5) * margin};
  R depth, dir[3], pos[3];
  if (!mpr_penetration(a, b, &depth, dir, pos)) return 0;
  if (dir[0] == R(0.0) && dir[1] == R(0.0) && dir[2] == R(0.0)) return 0;  // normal undefined
  con->dist = margin - depth;
  // MuJoCo's contact normal points from geom1 to geom2: libccd's direction is that of the separating translation of
  // obj1, i.e. v = s1 - s2 towards the origin's exit, which is geom1 -> geom2
  for (int k = 0; k < 3; k++) { con->frame[k] = dir[k]; con->pos[k] = pos[k]; }
  for (int k = 3; k < 9; k++) con->frame[k] = R(0.0);
  return 1;
}

// mjc_PlaneCylinder (engine_collision_primitive.c): nearest rim point of each cap, then two more points of the nearer
// cap that make a triangle with the first; at most 4 contacts
template <class R>
int plane_cylinder(Contact<R>* con, R margin, const R* pos1, const R* mat1, const R* pos2, const R* mat2, const double* size2) {
  using std::sqrt;
  R normal[3] = {mat1[2], mat1[5], mat1[8]};
  R axis[3] = {mat2[2], mat2[5], mat2[8]};
  R prjaxis = dot3(normal, axis);
  if (prjaxis > R(0.0)) { for (int k = 0; k < 3; k++) axis[k] = -axis[k]; prjaxis = -prjaxis; }
  R vec[3] = {pos2[0] - pos1[0], pos2[1] - pos1[1], pos2[2] - pos1[2]};
  R dist0 = dot3(vec, normal);
  for (int k = 0; k < 3; k++) vec[k] = axis[k] * prjaxis - normal[k];
  R len_sqr = dot3(vec, vec);
  if (len_sqr >= R(kMinVal) * R(kMinVal)) {
    R scl = R(size2[0]) / sqrt(len_sqr);
    for (int k = 0; k < 3; k++) vec[k] *= scl;
  } else {
    vec[0] = mat2[0] * R(size2[0]); vec[1] = mat2[3] * R(size2[0]); vec[2] = mat2[6] * R(size2[0]);
  }
  R prjvec = dot3(vec, normal);
  for (int k = 0; k < 3; k++) axis[k] *= R(size2[1]);
  prjaxis *= R(size2[1]);
  int cnt = 0;
  auto put = [&](R d, double sa, double sv, const R* side, double ss) {
    con[cnt].dist = d;
    for (int k = 0; k < 3; k++) {
      R p = pos2[k] + R(sv) * vec[k] + R(sa) * axis[k];
      if (side) p = p + R(ss) * side[k];
      con[cnt].pos[k] = p - normal[k] * d * R(0.5);
      con[cnt].frame[k] = normal[k];
    }
    for (int k = 3; k < 9; k++) con[cnt].frame[k] = R(0.0);
    cnt++;
  };
  if (dist0 + prjaxis + prjvec <= margin) put(dist0 + prjaxis + prjvec, 1.0, 1.0, nullptr, 0.0);
  else return 0;
  if (dist0 - prjaxis + prjvec <= margin) put(dist0 - prjaxis + prjvec, -1.0, 1.0, nullptr, 0.0);
  R prjvec1 = -prjvec * R(0.5);
  if (dist0 + prjaxis + prjvec1 <= margin) {
    R vec1[3];
    cross(vec1, vec, axis);
    normalize3(vec1);
    R s = R(size2[0]) * sqrt(R(3.0)) * R(0.5);
    for (int k = 0; k < 3; k++) vec1[k] *= s;
    put(dist0 + prjaxis + prjvec1, 1.0, -0.5, vec1, 1.0);
    put(dist0 + prjaxis + prjvec1, 1.0, -0.5, vec1, -1.0);
  }
  return cnt;
}

// mjc_PlaneConvex, deepest point only (see the header note)
template <class R>
int plane_convex(const Model* m, Contact<R>* con, R margin, const R* pos1, const R* mat1, int g2, const R* pos2, const R* mat2) {
  R normal[3] = {mat1[2], mat1[5], mat1[8]}, nd[3] = {-mat1[2], -mat1[5], -mat1[8]}, vec[3], dif[3];
  ConvexObj<R> o{m, g2, m->geom_type[g2], pos2, mat2, R(0.0)};
  convex_support(o, nd, vec);
  for (int k = 0; k < 3; k++) dif[k] = vec[k] - pos1[k];
  R dist = dot3(dif, normal);
  if (dist > margin) return 0;
  con->dist = dist;
  for (int k = 0; k < 3; k++) { con->pos[k] = vec[k] - R(0.5) * dist * normal[k]; con->frame[k] = normal[k]; }
  for (int k = 3; k < 9; k++) con->frame[k] = R(0.0);
  return 1;
}

}  // namespace ok
