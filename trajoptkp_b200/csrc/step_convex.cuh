// General convex collider of the batched step: mesh (convex hull) / box / sphere / capsule / cylinder pairs and
// plane-mesh, for the Franka scenes (reaching, box_sweep, push_*).
//
// MuJoCo: engine_collision_convex.c (mjc_Convex, mjc_PlaneConvex, mjccd_support / mjccd_center) on top of libccd's
// ccdMPRPenetration - Minkowski Portal Refinement (G. Snethen, XenoCollide, 2008) with tolerance 1e-6, at most 50
// refinement steps, each geom inflated by margin / 2; contact = {dist = margin - depth, normal = dir (geom1 ->
// geom2), pos = midpoint of the witnesses}; one contact per pair (mjENBL_MULTICCD is off by default).  Call sites on
// the reference path: every mj_step / mj_stepSkip (src/Differentiator/Differentiator.cpp:69,118; iLQR.cpp:222,914).
//
// The lanes of a group first filter the pairs of their world one pair per lane (bounding spheres, oriented boxes);
// the pairs that survive go through the narrow phase one after the other with the WHOLE group on each: the MPR state
// is replicated in every lane (group-uniform control flow, like the rest of the step) and the support function of a
// mesh - a scan of its merged vertices, whose maximiser is the hull's - is split over the lanes and reduced by
// shuffles (first maximal index wins, so the result equals the serial scan's).  A lane-per-pair narrow phase spent 64 %
// of the FD kernel's instructions in these scans at 1.8 active lanes (profiles/r1t_*).  plane-mesh yields the deepest
// vertex only (MuJoCo adds up to three hull neighbours; no hull graph is built).
#pragma once
#include "step.cuh"

namespace tk {

constexpr double kMprTol = 1e-6;
constexpr int kMprIter = 50;
constexpr double kCcdEps = 2.220446049250313e-16;

TK_DEV bool ccd_zero(double x) { return fabs(x) < kCcdEps; }
TK_DEV bool ccd_zero3(const double* v) { return ccd_zero(v[0]) && ccd_zero(v[1]) && ccd_zero(v[2]); }
TK_DEV bool is_convex_type(int t) {
  return t == GEOM_SPHERE || t == GEOM_CAPSULE || t == GEOM_CYLINDER || t == GEOM_BOX || t == GEOM_MESH;
}

struct CvxObj {
  int geom, type;
  const double* pos;
  const double* mat;
  double margin;
};

// index of the first vertex maximising v . ld, the n vertices split over the lanes of the group
template <int G, int NV>
TK_DEV int mesh_argmax(const Grp<G, NV>& g, const double* v, int n, const double* ld) {
  double bd = -1e300;
  int best = 0x7fffffff;
  for (int i = g.gl; i < n; i += G) {
    double dd = v[3 * i] * ld[0] + v[3 * i + 1] * ld[1] + v[3 * i + 2] * ld[2];
    if (dd > bd) { bd = dd; best = i; }
  }
#pragma unroll
  for (int off = G / 2; off > 0; off >>= 1) {
    double od = __shfl_xor_sync(g.mask, bd, off, G);
    int ob = __shfl_xor_sync(g.mask, best, off, G);
    if (od > bd || (od == bd && ob < best)) { bd = od; best = ob; }
  }
  return best;
}

// mjccd_support: farthest point of the inflated geom along the unit world direction dir (group-cooperative: every
// lane of the group calls it with the same arguments and receives the same point)
template <int G, int NV>
TK_DEV void convex_support(const DM& m, const Grp<G, NV>& g, const CvxObj& o, const double* dir, double* res) {
  const double* size = &m.geom_size[3 * o.geom];
  double ld[3], lp[3] = {0.0, 0.0, 0.0};
  for (int k = 0; k < 3; k++) ld[k] = o.mat[k] * dir[0] + o.mat[3 + k] * dir[1] + o.mat[6 + k] * dir[2];
  if (o.type == GEOM_MESH) {
    const double* v = &m.mesh_vert[3 * m.geom_meshadr[o.geom]];
    int best = mesh_argmax(g, v, m.geom_meshnum[o.geom], ld);
    for (int k = 0; k < 3; k++) lp[k] = v[3 * best + k];
  } else if (o.type == GEOM_BOX) {
    for (int k = 0; k < 3; k++) lp[k] = ld[k] >= 0.0 ? size[k] : -size[k];
  } else if (o.type == GEOM_SPHERE) {
    for (int k = 0; k < 3; k++) lp[k] = ld[k] * size[0];
  } else if (o.type == GEOM_CAPSULE) {
    for (int k = 0; k < 3; k++) lp[k] = ld[k] * size[0];
    lp[2] = lp[2] + (ld[2] >= 0.0 ? size[1] : -size[1]);
  } else if (o.type == GEOM_CYLINDER) {
    double r2 = ld[0] * ld[0] + ld[1] * ld[1];
    if (r2 > kMinVal * kMinVal) {
      double s = size[0] / sqrt(r2);
      lp[0] = ld[0] * s; lp[1] = ld[1] * s;
    }
    lp[2] = ld[2] >= 0.0 ? size[1] : -size[1];
  }
  for (int k = 0; k < 3; k++)
    res[k] = o.pos[k] + o.mat[3 * k] * lp[0] + o.mat[3 * k + 1] * lp[1] + o.mat[3 * k + 2] * lp[2] + dir[k] * o.margin;
}

struct MprPt { double v[3], v1[3], v2[3]; };

template <int G, int NV>
TK_DEV void mpr_support(const DM& m, const Grp<G, NV>& g, const CvxObj& a, const CvxObj& b, const double* dir, MprPt& p) {
  double nd[3] = {-dir[0], -dir[1], -dir[2]};
  convex_support(m, g, a, dir, p.v1);
  convex_support(m, g, b, nd, p.v2);
  for (int k = 0; k < 3; k++) p.v[k] = p.v1[k] - p.v2[k];
}

// closest point of triangle (a, b, c) to the origin (Ericson, Real-Time Collision Detection, 5.1.5)
TK_DEV void closest_on_triangle_to_origin(const double* a, const double* b, const double* c, double* out) {
  double ab[3], ac[3], ap[3], bp[3], cp[3];
  for (int k = 0; k < 3; k++) { ab[k] = b[k] - a[k]; ac[k] = c[k] - a[k]; ap[k] = -a[k]; bp[k] = -b[k]; cp[k] = -c[k]; }
  double d1 = dot3(ab, ap), d2 = dot3(ac, ap);
  if (d1 <= 0.0 && d2 <= 0.0) { for (int k = 0; k < 3; k++) out[k] = a[k]; return; }
  double d3 = dot3(ab, bp), d4 = dot3(ac, bp);
  if (d3 >= 0.0 && d4 <= d3) { for (int k = 0; k < 3; k++) out[k] = b[k]; return; }
  double vc = d1 * d4 - d3 * d2;
  if (vc <= 0.0 && d1 >= 0.0 && d3 <= 0.0) {
    double v = d1 / (d1 - d3);
    for (int k = 0; k < 3; k++) out[k] = a[k] + v * ab[k];
    return;
  }
  double d5 = dot3(ab, cp), d6 = dot3(ac, cp);
  if (d6 >= 0.0 && d5 <= d6) { for (int k = 0; k < 3; k++) out[k] = c[k]; return; }
  double vb = d5 * d2 - d1 * d6;
  if (vb <= 0.0 && d2 >= 0.0 && d6 <= 0.0) {
    double w = d2 / (d2 - d6);
    for (int k = 0; k < 3; k++) out[k] = a[k] + w * ac[k];
    return;
  }
  double va = d3 * d6 - d5 * d4;
  if (va <= 0.0 && (d4 - d3) >= 0.0 && (d5 - d6) >= 0.0) {
    double w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    for (int k = 0; k < 3; k++) out[k] = b[k] + w * (c[k] - b[k]);
    return;
  }
  double denom = 1.0 / (va + vb + vc);
  double v = vb * denom, w = vc * denom;
  for (int k = 0; k < 3; k++) out[k] = a[k] + ab[k] * v + ac[k] * w;
}

TK_DEV void mpr_portal_dir(const MprPt* p, double* dir) {
  double e1[3], e2[3];
  for (int k = 0; k < 3; k++) { e1[k] = p[2].v[k] - p[1].v[k]; e2[k] = p[3].v[k] - p[1].v[k]; }
  cross3(dir, e1, e2);
  normalize3(dir);
}

TK_DEV bool mpr_reach_tolerance(const MprPt* p, const MprPt& v4, const double* dir) {
  double dv4 = dot3(v4.v, dir);
  double d1 = dv4 - dot3(p[1].v, dir), d2 = dv4 - dot3(p[2].v, dir), d3 = dv4 - dot3(p[3].v, dir);
  double mn = d1 < d2 ? d1 : d2;
  if (d3 < mn) mn = d3;
  return mn <= kMprTol;
}

TK_DEV void mpr_expand(MprPt* p, const MprPt& v4) {
  double v4v0[3];
  cross3(v4v0, v4.v, p[0].v);
  if (dot3(p[1].v, v4v0) > 0.0) {
    if (dot3(p[2].v, v4v0) > 0.0) p[1] = v4; else p[3] = v4;
  } else {
    if (dot3(p[3].v, v4v0) > 0.0) p[2] = v4; else p[1] = v4;
  }
}

// ccdMPRPenetration: 1 and depth / dir / pos when the inflated geoms intersect, else 0
template <int G, int NV>
TK_DEV int mpr_penetration(const DM& m, const Grp<G, NV>& g, const CvxObj& a, const CvxObj& b, double* depth, double* dir_out,
                           double* pos_out) {
  MprPt p[4];  // p[0] = interior point, p[1..3] = portal
  double dir[3], va[3], vb[3];
  for (int k = 0; k < 3; k++) { p[0].v1[k] = a.pos[k]; p[0].v2[k] = b.pos[k]; p[0].v[k] = a.pos[k] - b.pos[k]; }
  if (ccd_zero3(p[0].v)) p[0].v[0] = p[0].v[0] + 10.0 * kCcdEps;
  for (int k = 0; k < 3; k++) dir[k] = -p[0].v[k];
  normalize3(dir);
  mpr_support(m, g, a, b, dir, p[1]);
  double d = dot3(p[1].v, dir);
  if (ccd_zero(d) || d < 0.0) return 0;
  cross3(dir, p[0].v, p[1].v);
  if (ccd_zero3(dir)) {
    if (ccd_zero3(p[1].v)) {
      *depth = 0.0;
      for (int k = 0; k < 3; k++) { dir_out[k] = 0.0; pos_out[k] = 0.5 * (p[1].v1[k] + p[1].v2[k]); }
      return 1;
    }
    for (int k = 0; k < 3; k++) { dir_out[k] = p[1].v[k]; pos_out[k] = 0.5 * (p[1].v1[k] + p[1].v2[k]); }
    double nn = sqrt(dot3(dir_out, dir_out));
    *depth = nn;
    for (int k = 0; k < 3; k++) dir_out[k] = dir_out[k] / nn;
    return 1;
  }
  normalize3(dir);
  mpr_support(m, g, a, b, dir, p[2]);
  d = dot3(p[2].v, dir);
  if (ccd_zero(d) || d < 0.0) return 0;
  for (int k = 0; k < 3; k++) { va[k] = p[1].v[k] - p[0].v[k]; vb[k] = p[2].v[k] - p[0].v[k]; }
  cross3(dir, va, vb);
  normalize3(dir);
  if (dot3(dir, p[0].v) > 0.0) {
    MprPt t = p[1]; p[1] = p[2]; p[2] = t;
    for (int k = 0; k < 3; k++) dir[k] = -dir[k];
  }
  for (int guard = 0;; guard++) {
    if (guard > 100) return 0;
    mpr_support(m, g, a, b, dir, p[3]);
    d = dot3(p[3].v, dir);
    if (ccd_zero(d) || d < 0.0) return 0;
    bool cont = false;
    cross3(va, p[1].v, p[3].v);
    d = dot3(va, p[0].v);
    if (d < 0.0 && !ccd_zero(d)) { p[2] = p[3]; cont = true; }
    if (!cont) {
      cross3(va, p[3].v, p[2].v);
      d = dot3(va, p[0].v);
      if (d < 0.0 && !ccd_zero(d)) { p[1] = p[3]; cont = true; }
    }
    if (!cont) break;
    for (int k = 0; k < 3; k++) { va[k] = p[1].v[k] - p[0].v[k]; vb[k] = p[2].v[k] - p[0].v[k]; }
    cross3(dir, va, vb);
    normalize3(dir);
  }
  MprPt v4;
  for (int guard = 0;; guard++) {
    mpr_portal_dir(p, dir);
    d = dot3(dir, p[1].v);
    if (ccd_zero(d) || d > 0.0) break;
    mpr_support(m, g, a, b, dir, v4);
    d = dot3(v4.v, dir);
    if (!(ccd_zero(d) || d > 0.0) || mpr_reach_tolerance(p, v4, dir) || guard > 1000) return 0;
    mpr_expand(p, v4);
  }
  for (int it = 0;; it++) {
    mpr_portal_dir(p, dir);
    mpr_support(m, g, a, b, dir, v4);
    if (mpr_reach_tolerance(p, v4, dir) || it > kMprIter) {
      double w[3];
      closest_on_triangle_to_origin(p[1].v, p[2].v, p[3].v, w);
      *depth = sqrt(dot3(w, w));
      if (ccd_zero3(w)) for (int k = 0; k < 3; k++) w[k] = dir[k];
      normalize3(w);
      for (int k = 0; k < 3; k++) dir_out[k] = w[k];
      double b4[4], t[3];
      cross3(t, p[1].v, p[2].v); b4[0] = dot3(t, p[3].v);
      cross3(t, p[3].v, p[2].v); b4[1] = dot3(t, p[0].v);
      cross3(t, p[0].v, p[1].v); b4[2] = dot3(t, p[3].v);
      cross3(t, p[2].v, p[1].v); b4[3] = dot3(t, p[0].v);
      double sum = b4[0] + b4[1] + b4[2] + b4[3];
      if (ccd_zero(sum) || sum < 0.0) {
        b4[0] = 0.0;
        cross3(t, p[2].v, p[3].v); b4[1] = dot3(t, dir);
        cross3(t, p[3].v, p[1].v); b4[2] = dot3(t, dir);
        cross3(t, p[1].v, p[2].v); b4[3] = dot3(t, dir);
        sum = b4[1] + b4[2] + b4[3];
      }
      double inv = 1.0 / sum;
      for (int k = 0; k < 3; k++) {
        double p1 = 0.0, p2 = 0.0;
        for (int i = 0; i < 4; i++) { p1 = p1 + b4[i] * p[i].v1[k]; p2 = p2 + b4[i] * p[i].v2[k]; }
        pos_out[k] = 0.5 * (p1 + p2) * inv;
      }
      return 1;
    }
    mpr_expand(p, v4);
  }
}

// Separating-axis test of the two geoms' oriented bounding boxes (half-extents h1 / h2 in the geom frames, inflated
// by margin / 2 each; mesh geoms carry the box of their vertices in geom_size).  true = the boxes are disjoint, so
// the convex shapes inside them are too and the narrow phase would return no contact: a cull that cannot change a
// result (MuJoCo's mj_collideGeoms applies the same kind of box filter before the narrow phase).
TK_DEV bool obb_disjoint(const double* pos1, const double* mat1, const double* h1, const double* pos2, const double* mat2,
                         const double* h2, double margin) {
  double a[3] = {h1[0] + 0.5 * margin, h1[1] + 0.5 * margin, h1[2] + 0.5 * margin};
  double b[3] = {h2[0] + 0.5 * margin, h2[1] + 0.5 * margin, h2[2] + 0.5 * margin};
  double R[9], AR[9], t[3];
  double d[3] = {pos2[0] - pos1[0], pos2[1] - pos1[1], pos2[2] - pos1[2]};
  for (int i = 0; i < 3; i++) {
    t[i] = mat1[i] * d[0] + mat1[3 + i] * d[1] + mat1[6 + i] * d[2];
    for (int j = 0; j < 3; j++) {
      double r = mat1[i] * mat2[j] + mat1[3 + i] * mat2[3 + j] + mat1[6 + i] * mat2[6 + j];
      R[3 * i + j] = r;
      AR[3 * i + j] = fabs(r) + 1e-9;  // slack: near-parallel edges make the cross-product axes degenerate
    }
  }
  for (int i = 0; i < 3; i++)
    if (fabs(t[i]) > a[i] + b[0] * AR[3 * i] + b[1] * AR[3 * i + 1] + b[2] * AR[3 * i + 2]) return true;
  for (int j = 0; j < 3; j++)
    if (fabs(t[0] * R[j] + t[1] * R[3 + j] + t[2] * R[6 + j]) > a[0] * AR[j] + a[1] * AR[3 + j] + a[2] * AR[6 + j] + b[j]) return true;
  for (int i = 0; i < 3; i++) {
    int i1 = (i + 1) % 3, i2 = (i + 2) % 3;
    for (int j = 0; j < 3; j++) {
      int j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      double ra = a[i1] * AR[3 * i2 + j] + a[i2] * AR[3 * i1 + j];
      double rb = b[j1] * AR[3 * i + j2] + b[j2] * AR[3 * i + j1];
      if (fabs(t[i2] * R[3 * i1 + j] - t[i1] * R[3 * i2 + j]) > ra + rb) return true;
    }
  }
  return false;
}

// pairs MuJoCo's collision table sends to mjc_Convex among the geom types in scope: everything with a mesh, and the
// cylinder pairs without a native collider (cylinder-cylinder, box-cylinder)
TK_DEV bool uses_convex_collider(int t1, int t2) {
  if (!is_convex_type(t1) || !is_convex_type(t2)) return false;
  if (t1 == GEOM_MESH || t2 == GEOM_MESH) return true;
  return (t1 == GEOM_CYLINDER && t2 == GEOM_CYLINDER) || (t1 == GEOM_CYLINDER && t2 == GEOM_BOX) ||
         (t1 == GEOM_BOX && t2 == GEOM_CYLINDER);
}

// mjc_PlaneCylinder (engine_collision_primitive.c): the rim point nearest the plane on each cap, plus two points of the
// nearer cap that complete a triangle; at most 4 contacts.  dist[], pos[3*], frame[6*] are slot arrays.
TK_DEV int plane_cylinder(double margin, const double* pos1, const double* mat1, const double* pos2, const double* mat2,
                          const double* size2, double* dist, double* pos, double* frame) {
  double normal[3] = {mat1[2], mat1[5], mat1[8]};
  double axis[3] = {mat2[2], mat2[5], mat2[8]};
  double prjaxis = dot3(normal, axis);
  if (prjaxis > 0.0) { for (int k = 0; k < 3; k++) axis[k] = -axis[k]; prjaxis = -prjaxis; }
  double vec[3] = {pos2[0] - pos1[0], pos2[1] - pos1[1], pos2[2] - pos1[2]};
  double dist0 = dot3(vec, normal);
  for (int k = 0; k < 3; k++) vec[k] = axis[k] * prjaxis - normal[k];
  double len_sqr = dot3(vec, vec);
  if (len_sqr >= kMinVal * kMinVal) {
    double scl = size2[0] / sqrt(len_sqr);
    for (int k = 0; k < 3; k++) vec[k] *= scl;
  } else {  // disc parallel to the plane: any radial direction, MuJoCo takes the cylinder's x axis
    vec[0] = mat2[0] * size2[0]; vec[1] = mat2[3] * size2[0]; vec[2] = mat2[6] * size2[0];
  }
  double prjvec = dot3(vec, normal);
  for (int k = 0; k < 3; k++) axis[k] *= size2[1];
  prjaxis *= size2[1];
  int cnt = 0;
  auto put = [&](double d, double sa, double sv, const double* side, double ss) {
    dist[cnt] = d;
    for (int k = 0; k < 3; k++) {
      pos[3 * cnt + k] = pos2[k] + sv * vec[k] + sa * axis[k] + (side ? ss * side[k] : 0.0) - normal[k] * d * 0.5;
      frame[6 * cnt + k] = normal[k];
      frame[6 * cnt + 3 + k] = 0.0;
    }
    cnt++;
  };
  if (dist0 + prjaxis + prjvec <= margin) put(dist0 + prjaxis + prjvec, 1.0, 1.0, nullptr, 0.0);
  else return 0;
  if (dist0 - prjaxis + prjvec <= margin) put(dist0 - prjaxis + prjvec, -1.0, 1.0, nullptr, 0.0);
  double prjvec1 = -prjvec * 0.5;
  if (dist0 + prjaxis + prjvec1 <= margin) {
    double vec1[3];
    cross3(vec1, vec, axis);
    normalize3(vec1);
    double s = size2[0] * sqrt(3.0) * 0.5;
    for (int k = 0; k < 3; k++) vec1[k] *= s;
    put(dist0 + prjaxis + prjvec1, 1.0, -0.5, vec1, 1.0);
    put(dist0 + prjaxis + prjvec1, 1.0, -0.5, vec1, -1.0);
  }
  return cnt;
}

// mjc_Convex: writes dist, pos[3], frame[0:6] (tangent hint zero); returns the number of contacts (0 or 1)
template <int G, int NV>
TK_DEV int convex_convex(const DM& m, const Grp<G, NV>& g, double margin, int g1, const double* pos1, const double* mat1, int g2, const double* pos2,
                         const double* mat2, double* dist, double* pos, double* frame) {
  CvxObj a{g1, m.geom_type[g1], pos1, mat1, 0.5 * margin};
  CvxObj b{g2, m.geom_type[g2], pos2, mat2, 0.5 * margin};
  double depth, dir[3], cp[3];
  if (!mpr_penetration(m, g, a, b, &depth, dir, cp)) return 0;
  if (dir[0] == 0.0 && dir[1] == 0.0 && dir[2] == 0.0) return 0;
  *dist = margin - depth;
  for (int k = 0; k < 3; k++) { frame[k] = dir[k]; pos[k] = cp[k]; frame[3 + k] = 0.0; }
  return 1;
}

// mjc_PlaneConvex, deepest point only
template <int G, int NV>
TK_DEV int plane_convex(const DM& m, const Grp<G, NV>& g, double margin, const double* pos1, const double* mat1, int g2, const double* pos2,
                        const double* mat2, double* dist, double* pos, double* frame) {
  double nrm[3] = {mat1[2], mat1[5], mat1[8]}, nd[3] = {-mat1[2], -mat1[5], -mat1[8]}, vec[3], dif[3];
  CvxObj o{g2, m.geom_type[g2], pos2, mat2, 0.0};
  convex_support(m, g, o, nd, vec);
  for (int k = 0; k < 3; k++) dif[k] = vec[k] - pos1[k];
  double dd = dot3(dif, nrm);
  if (dd > margin) return 0;
  *dist = dd;
  for (int k = 0; k < 3; k++) { pos[k] = vec[k] - 0.5 * dd * nrm[k]; frame[k] = nrm[k]; frame[3 + k] = 0.0; }
  return 1;
}

}  // namespace tk
