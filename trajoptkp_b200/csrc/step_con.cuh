// Collision + constraint construction of the batched step (MuJoCo: mj_collision, mj_makeConstraint,
// mj_referenceConstraint) for joint limits and frictionless / pyramidal contacts.
#pragma once
#include "step_pos.cuh"
#include "step_convex.cuh"

namespace tk {

struct ConParam {
  double margin, gap, mu, solref[2], solimp[5];
  int dim;
};

// mj_contactParam for a geom pair
TK_DEV void contact_param(const DM& m, int g1, int g2, ConParam& c) {
  c.margin = fmax(m.geom_margin[g1], m.geom_margin[g2]);
  c.gap = fmax(m.geom_gap[g1], m.geom_gap[g2]);
  int p1 = m.geom_priority[g1], p2 = m.geom_priority[g2];
  if (p1 != p2) {
    int gg = p1 > p2 ? g1 : g2;
    c.dim = m.geom_condim[gg];
    for (int k = 0; k < 2; k++) c.solref[k] = m.geom_solref[2 * gg + k];
    for (int k = 0; k < 5; k++) c.solimp[k] = m.geom_solimp[5 * gg + k];
    c.mu = m.geom_friction[3 * gg];
  } else {
    c.dim = max(m.geom_condim[g1], m.geom_condim[g2]);
    double s1 = m.geom_solmix[g1], s2 = m.geom_solmix[g2], mix;
    if (s1 >= kMinVal && s2 >= kMinVal) mix = s1 / (s1 + s2);
    else if (s1 < kMinVal && s2 < kMinVal) mix = 0.5;
    else if (s1 < kMinVal) mix = 0.0;
    else mix = 1.0;
    double a0 = m.geom_solref[2 * g1], a1 = m.geom_solref[2 * g1 + 1], b0 = m.geom_solref[2 * g2], b1 = m.geom_solref[2 * g2 + 1];
    if (a0 > 0 && b0 > 0) { c.solref[0] = mix * a0 + (1 - mix) * b0; c.solref[1] = mix * a1 + (1 - mix) * b1; }
    else { c.solref[0] = fmin(a0, b0); c.solref[1] = fmin(a1, b1); }
    for (int k = 0; k < 5; k++) c.solimp[k] = mix * m.geom_solimp[5 * g1 + k] + (1 - mix) * m.geom_solimp[5 * g2 + k];
    c.mu = fmax(m.geom_friction[3 * g1], m.geom_friction[3 * g2]);
  }
}

TK_DEV int pair_slots(int t1, int t2) {
  if (t1 == GEOM_PLANE) return (t2 == GEOM_SPHERE || t2 == GEOM_MESH) ? 1 : (t2 == GEOM_CAPSULE ? 2 : 4);
  if (t1 == GEOM_BOX && t2 == GEOM_BOX) return 8;
  if (t1 == GEOM_CAPSULE && t2 == GEOM_BOX) return 2;
  return 1;
}

// row table entry
TK_DEV int row_pack(int type, int id, int sub) { return type | ((sub + 1) << 2) | (id << 5); }
TK_DEV int row_type_of(int info) { return info & 3; }
TK_DEV int row_sub_of(int info) { return ((info >> 2) & 7) - 1; }
TK_DEV int row_id_of(int info) { return info >> 5; }

// complete frame[0:3] (normal) and the tangent hint frame[3:6] to an orthonormal pair (mju_makeFrame); the third axis
// is cross(normal, tangent1), formed where it is used
TK_DEV void make_frame(double* f) {
  normalize3(f);
  double yn = sqrt(dot3(f + 3, f + 3));
  if (yn < 0.5) {
    f[3] = 0.0; f[4] = 0.0; f[5] = 0.0;
    if (f[1] < 0.5 && f[1] > -0.5) f[4] = 1.0; else f[5] = 1.0;
  }
  double t = dot3(f, f + 3);
  f[3] -= t * f[0]; f[4] -= t * f[1]; f[5] -= t * f[2];
  normalize3(f + 3);
}

// plane vs sphere of given centre/radius; writes dist, pos, frame[0:9] (tangents zero); returns 1 on contact
TK_DEV int plane_sphere(double margin, const double* ppos, const double* pmat, const double* c, double radius, double* dist,
                        double* pos, double* frame) {
  double nrm[3] = {pmat[2], pmat[5], pmat[8]};
  double d[3] = {c[0] - ppos[0], c[1] - ppos[1], c[2] - ppos[2]};
  double cd = dot3(d, nrm);
  if (cd > margin + radius) return 0;
  *dist = cd - radius;
  double s = -(*dist) * 0.5 - radius;
  for (int k = 0; k < 3; k++) { pos[k] = c[k] + nrm[k] * s; frame[k] = nrm[k]; }
  for (int k = 3; k < 6; k++) frame[k] = 0.0;
  return 1;
}

// mj_collision over the static pair list: lanes = pairs, fixed slots per pair, then ordered compaction
template <int G, int NV>
TK_DEV void collision(const DM& m, WS& w, const Grp<G, NV>& g, int& status) {
  // slot offsets are the prefix of pair_slots: recomputed on the fly (npair is small)
  for (int p = g.gl; p < m.npair; p += G) {
    int g1 = m.pair_geom1[p], g2 = m.pair_geom2[p];
    int t1 = m.geom_type[g1], t2 = m.geom_type[g2];
    int slot = m.pair_slotadr[p];  // prefix of pair_slots(), laid down by the model compiler
    double margin = fmax(m.geom_margin[g1], m.geom_margin[g2]);
    double pos1[3], mat1[9], pos2[3], mat2[9];
    geom_pose(m, w, g1, pos1, mat1);
    geom_pose(m, w, g2, pos2, mat2);
    int n = 0;
    if (t1 == GEOM_PLANE) {
      double nrm[3] = {mat1[2], mat1[5], mat1[8]};
      double d[3] = {pos2[0] - pos1[0], pos2[1] - pos1[1], pos2[2] - pos1[2]};
      if (dot3(d, nrm) <= margin + m.geom_rbound[g2]) {
        if (t2 == GEOM_SPHERE) {
          n = plane_sphere(margin, pos1, mat1, pos2, m.geom_size[3 * g2], &w.con_dist[slot], &w.con_pos[3 * slot], &w.con_frame[6 * slot]);
        } else if (t2 == GEOM_CAPSULE) {
          double axis[3] = {mat2[2], mat2[5], mat2[8]}, c[3];
          double hl = m.geom_size[3 * g2 + 1], r = m.geom_size[3 * g2];
          for (int k = 0; k < 3; k++) c[k] = pos2[k] + axis[k] * hl;
          int n1 = plane_sphere(margin, pos1, mat1, c, r, &w.con_dist[slot], &w.con_pos[3 * slot], &w.con_frame[6 * slot]);
          for (int k = 0; k < 3; k++) c[k] = pos2[k] - axis[k] * hl;
          int n2 = plane_sphere(margin, pos1, mat1, c, r, &w.con_dist[slot + n1], &w.con_pos[3 * (slot + n1)], &w.con_frame[6 * (slot + n1)]);
          n = n1 + n2;
          for (int i = 0; i < n; i++)
            for (int k = 0; k < 3; k++) w.con_frame[6 * (slot + i) + 3 + k] = axis[k];
        } else if (t2 == GEOM_BOX) {
          double nrm2[3] = {mat1[2], mat1[5], mat1[8]};
          double dist0 = dot3(d, nrm2);
          for (int i = 0; i < 8 && n < 4; i++) {
            double vec[3], corner[3];
            vec[0] = (i & 1) ? m.geom_size[3 * g2] : -m.geom_size[3 * g2];
            vec[1] = (i & 2) ? m.geom_size[3 * g2 + 1] : -m.geom_size[3 * g2 + 1];
            vec[2] = (i & 4) ? m.geom_size[3 * g2 + 2] : -m.geom_size[3 * g2 + 2];
            mat_vec3(corner, mat2, vec);
            double ld = dot3(nrm2, corner);
            if (dist0 + ld > margin || ld > 0.0) continue;
            double cd = dist0 + ld;
            int s = slot + n;
            w.con_dist[s] = cd;
            for (int k = 0; k < 3; k++) { w.con_pos[3 * s + k] = corner[k] + pos2[k] - nrm2[k] * cd * 0.5; w.con_frame[6 * s + k] = nrm2[k]; }
            for (int k = 3; k < 6; k++) w.con_frame[6 * s + k] = 0.0;
            n++;
          }
        } else if (NV == 0 && t2 == GEOM_CYLINDER) {
          n = plane_cylinder(margin, pos1, mat1, pos2, mat2, &m.geom_size[3 * g2], &w.con_dist[slot], &w.con_pos[3 * slot],
                             &w.con_frame[6 * slot]);
        } else if (NV == 0 && t2 == GEOM_MESH) {
          n = -1;  // narrow phase by the whole group, below
        } else {
          status |= ST_UNSUPPORTED;
        }
      }
    } else if (NV == 0 && uses_convex_collider(t1, t2)) {
      // mj_collideGeoms: bounding-sphere filter, then the general convex collider.  Only the generic (NV = 0)
      // instantiation carries this code: the specialised walker / acrobot kernels have no mesh geoms
      double d[3] = {pos2[0] - pos1[0], pos2[1] - pos1[1], pos2[2] - pos1[2]};
      double bound = m.geom_rbound[g1] + m.geom_rbound[g2] + margin;
      bool boxed = (t1 == GEOM_MESH || t1 == GEOM_BOX) && (t2 == GEOM_MESH || t2 == GEOM_BOX);
      if (dot3(d, d) <= bound * bound &&
          !(boxed && obb_disjoint(pos1, mat1, &m.geom_size[3 * g1], pos2, mat2, &m.geom_size[3 * g2], margin)))
        n = -1;  // narrow phase by the whole group, below
    } else {
      status |= ST_UNSUPPORTED;
    }
    for (int i = 0; i < n; i++) make_frame(&w.con_frame[6 * (slot + i)]);
    w.cnt[p] = n;
  }
  g.sync();
  TK_CLK(w, g, 16);
  if (NV == 0) {
    // convex narrow phase: the pairs that passed the filters, one after the other, all lanes of the group on each
    for (int p = 0; p < m.npair; p++) {
      if (w.cnt[p] < 0) {  // group-uniform: every lane reads the same flag
        TK_CNT(w, g, 23);
        int g1 = m.pair_geom1[p], g2 = m.pair_geom2[p];
        int t1 = m.geom_type[g1];
        int slot = m.pair_slotadr[p];
        double margin = fmax(m.geom_margin[g1], m.geom_margin[g2]);
        double pos1[3], mat1[9], pos2[3], mat2[9], dist, cpos[3], frame[6];
        geom_pose(m, w, g1, pos1, mat1);
        geom_pose(m, w, g2, pos2, mat2);
        int n = t1 == GEOM_PLANE ? plane_convex(m, g, margin, pos1, mat1, g2, pos2, mat2, &dist, cpos, frame)
                                 : convex_convex(m, g, margin, g1, pos1, mat1, g2, pos2, mat2, &dist, cpos, frame);
        g.sync();
        if (g.gl == 0) {
          if (n) {
            make_frame(frame);
            w.con_dist[slot] = dist;
            for (int k = 0; k < 3; k++) w.con_pos[3 * slot + k] = cpos[k];
            for (int k = 0; k < 6; k++) w.con_frame[6 * slot + k] = frame[k];
          }
          w.cnt[p] = n;
        }
        g.sync();
      }
    }
  }
  TK_CLK(w, g, 17);
  // ordered compaction (pair order, then contact order), done by one lane: contact counts are tiny
  if (g.gl == 0) {
    int dst = 0;
    for (int p = 0; p < m.npair; p++) {
      int n = w.cnt[p], slot = m.pair_slotadr[p];
      for (int i = 0; i < n; i++) {
        int s = slot + i;
        if (s != dst) {
          w.con_dist[dst] = w.con_dist[s];
          for (int k = 0; k < 3; k++) w.con_pos[3 * dst + k] = w.con_pos[3 * s + k];
          for (int k = 0; k < 6; k++) w.con_frame[6 * dst + k] = w.con_frame[6 * s + k];
        }
        w.con_pair[dst] = p;
        dst++;
      }
    }
    w.cnt[m.npair > m.njnt ? m.npair : m.njnt] = dst;
  }
  g.sync();
  w.ncon = w.cnt[m.npair > m.njnt ? m.npair : m.njnt];
  g.sync();
}

// impedance + stiffness/damping for one row (getimpedance, mj_makeImpedance)
TK_DEV void make_kbip(const DM& m, const double* solref_in, const double* solimp_in, double pos, double margin, double& K,
                      double& B, double& imp) {
  double dmin = fmin(fmax(solimp_in[0], kMinImp), kMaxImp), dmax = fmin(fmax(solimp_in[1], kMinImp), kMaxImp);
  double width = fmax(solimp_in[2], 0.0), mid = fmin(fmax(solimp_in[3], kMinImp), kMaxImp), power = fmax(solimp_in[4], 1.0);
  double tc = solref_in[0], dr = solref_in[1];
  if (tc > 0.0 && tc < 2.0 * m.timestep) tc = 2.0 * m.timestep;
  double x = pos - margin;
  if (!(dmin != dmax) || width < kMinVal) {
    imp = 0.5 * (dmin + dmax);
  } else {
    x = fabs(x / width);
    if (x >= 1.0 || x <= 0.0) {
      imp = (x >= 1.0) ? dmax : dmin;
    } else {
      double y;
      // power 2 (the default solimp) needs no pow(): pow(a, 1) = a and pow(a, 2) = a * a, correctly rounded
      if (power == 1.0) y = x;
      else if (power == 2.0) y = (x <= mid) ? (1.0 / mid) * (x * x) : 1.0 - (1.0 / (1.0 - mid)) * ((1.0 - x) * (1.0 - x));
      else if (x <= mid) y = (1.0 / pow(mid, power - 1.0)) * pow(x, power);
      else y = 1.0 - (1.0 / pow(1.0 - mid, power - 1.0)) * pow(1.0 - x, power);
      imp = dmin + y * (dmax - dmin);
    }
  }
  if (tc > 0.0) { K = 1.0 / (dmax * dmax * tc * tc * dr * dr); B = 2.0 / (dmax * tc); }
  else { K = -tc / (dmax * dmax); B = -dr / dmax; }
}

// mj_makeConstraint: row table (limits then contacts), Jacobian rows, regularisation
template <int G, int NV>
TK_DEV void make_constraint(const DM& m, WS& w, const Grp<G, NV>& g, int& status) {
  const int nv = NV ? NV : m.nv;
  const int cnt_slot = (m.npair > m.njnt ? m.npair : m.njnt) + 1;
  // 1. row table: limit rows in joint order (lower side first), then 1 or 4 rows per included contact in contact order.
  // Lanes take one joint / one contact each and find their first row from ballots over the group; w.con_pair[c] gets
  // the contact's first row in its upper half (0 = no rows).  Counts that do not fit the row capacity are rare (the
  // world is then flagged and, below full capacity, re-run): one lane redoes the table in sequence with the
  // reference's skip rules.
  const unsigned below = (1u << g.gl) - 1u;
  int r = 0;
  bool fits = true;
  for (int j0 = 0; j0 < m.njnt; j0 += G) {
    const int j = j0 + g.gl;
    int bits = 0;
    double dlo = 0.0, dhi = 0.0, margin = 0.0;
    if (j < m.njnt && m.jnt_limited[j]) {
      const int jt = m.jnt_type[j];
      if (jt == JNT_HINGE || jt == JNT_SLIDE) {
        const double value = w.qpos[m.jnt_qposadr[j]];
        margin = m.jnt_margin[j];
        dlo = -1 * (m.jnt_range[2 * j] - value);
        dhi = 1 * (m.jnt_range[2 * j + 1] - value);
        bits = (dlo < margin ? 1 : 0) | (dhi < margin ? 2 : 0);
      }
    }
    const unsigned b0 = g.ballot(bits & 1), b1 = g.ballot(bits & 2);
    int mine = r + __popc(b0 & below) + __popc(b1 & below);
    r += __popc(b0) + __popc(b1);
    if (r <= w.ecap) {
      if (bits & 1) { w.row_info[mine] = row_pack(ROW_LIMIT, j, -1); w.efc_pos[mine] = dlo; w.efc_margin[mine] = margin; mine++; }
      if (bits & 2) { w.row_info[mine] = row_pack(ROW_LIMIT, j, 1); w.efc_pos[mine] = dhi; w.efc_margin[mine] = margin; }
    } else {
      fits = false;
    }
  }
  for (int c0 = 0; c0 < w.ncon && fits; c0 += G) {
    const int c = c0 + g.gl;
    int rows = 0, pr = 0;
    double inc = 0.0;
    if (c < w.ncon) {
      pr = w.con_pair[c];
      ConParam cp;
      contact_param(m, m.pair_geom1[pr], m.pair_geom2[pr], cp);
      inc = cp.margin - cp.gap;
      if (w.con_dist[c] < inc) {  // otherwise excluded (in the gap)
        rows = cp.dim == 1 ? 1 : (cp.dim == 3 ? 4 : 0);
        if (rows == 0) status |= ST_UNSUPPORTED;
      }
    }
    const unsigned b1 = g.ballot(rows == 1), b4 = g.ballot(rows == 4);
    const int mine = r + __popc(b1 & below) + 4 * __popc(b4 & below);
    r += __popc(b1) + 4 * __popc(b4);
    if (r <= w.ecap) {
      if (c < w.ncon) w.con_pair[c] = pr | ((rows ? mine + 1 : 0) << 16);
      for (int k = 0; k < rows; k++) {
        w.row_info[mine + k] = row_pack(rows == 1 ? ROW_FRICTIONLESS : ROW_PYRAMIDAL, c, k);
        w.efc_pos[mine + k] = w.con_dist[c]; w.efc_margin[mine + k] = inc;
      }
    } else {
      fits = false;
    }
  }
  g.sync();
  if (!fits) {
    if (g.gl == 0) {
      r = 0;
      for (int j = 0; j < m.njnt; j++) {
        if (!m.jnt_limited[j]) continue;
        int jt = m.jnt_type[j];
        if (jt != JNT_HINGE && jt != JNT_SLIDE) continue;
        double value = w.qpos[m.jnt_qposadr[j]], margin = m.jnt_margin[j];
        for (int side = -1; side <= 1; side += 2) {
          double dist = side * (m.jnt_range[2 * j + (side + 1) / 2] - value);
          if (dist < margin) {
            if (r >= w.ecap) { status |= ST_OVERFLOW; continue; }
            w.row_info[r] = row_pack(ROW_LIMIT, j, side);
            w.efc_pos[r] = dist; w.efc_margin[r] = margin;
            r++;
          }
        }
      }
      for (int c = 0; c < w.ncon; c++) {
        int pr = w.con_pair[c] & 0xffff;
        w.con_pair[c] = pr;
        ConParam cp;
        contact_param(m, m.pair_geom1[pr], m.pair_geom2[pr], cp);
        double inc = cp.margin - cp.gap;
        if (w.con_dist[c] >= inc) continue;  // excluded (in the gap)
        int rows = cp.dim == 1 ? 1 : (cp.dim == 3 ? 4 : -1);
        if (rows < 0) { status |= ST_UNSUPPORTED; continue; }
        if (r + rows > w.ecap) { status |= ST_OVERFLOW; continue; }
        w.con_pair[c] = pr | ((r + 1) << 16);
        for (int k = 0; k < rows; k++) {
          w.row_info[r] = row_pack(rows == 1 ? ROW_FRICTIONLESS : ROW_PYRAMIDAL, c, k);
          w.efc_pos[r] = w.con_dist[c]; w.efc_margin[r] = inc;
          r++;
        }
      }
      w.cnt[cnt_slot] = r;
    }
    g.sync();
    r = w.cnt[cnt_slot];
  }
  const int nefc = r;
  w.nefc = nefc;
  // 2. Jacobian: limit rows are +-1 at the joint's dof; a contact's rows share the difference of the two bodies' point
  // Jacobians, computed once per (contact, dof) (mj_makeConstraint builds the pyramid edges from one contact Jacobian)
  for (int e = g.gl; e < nefc * nv; e += G) {
    const int rr = e / nv, i = e - rr * nv;
    const int info = w.row_info[rr];
    if (row_type_of(info) != ROW_LIMIT) break;  // limit rows come first
    w.J[e] = (m.jnt_dofadr[row_id_of(info)] == i) ? (double)(-row_sub_of(info)) : 0.0;
  }
  for (int e = g.gl; e < w.ncon * nv; e += G) {
    const int c = e / nv, i = e - c * nv;
    const int packed = w.con_pair[c], row0 = (packed >> 16) - 1;
    if (row0 < 0) continue;
    const int pr = packed & 0xffff;
    const int g1 = m.pair_geom1[pr], g2 = m.pair_geom2[pr];
    double c1[3], c2[3], d[3];
    jac_col(m, w, c1, &w.con_pos[3 * c], m.geom_bodyid[g1], i);
    jac_col(m, w, c2, &w.con_pos[3 * c], m.geom_bodyid[g2], i);
    d[0] = c2[0] - c1[0]; d[1] = c2[1] - c1[1]; d[2] = c2[2] - c1[2];
    const double* f = &w.con_frame[6 * c];
    const double jn = f[0] * d[0] + f[1] * d[1] + f[2] * d[2];
    if (row_type_of(w.row_info[row0]) == ROW_FRICTIONLESS) {
      w.J[row0 * nv + i] = jn;
    } else {
      double t2[3];
      cross3(t2, f, f + 3);
      const double jt1 = f[3] * d[0] + f[4] * d[1] + f[5] * d[2];
      const double jt2 = t2[0] * d[0] + t2[1] * d[1] + t2[2] * d[2];
      double mu = fmax(m.geom_friction[3 * g1], m.geom_friction[3 * g2]);
      if (m.geom_priority[g1] != m.geom_priority[g2])
        mu = m.geom_friction[3 * (m.geom_priority[g1] > m.geom_priority[g2] ? g1 : g2)];
      w.J[row0 * nv + i] = jn + mu * jt1;        // rows: +t1, -t1, +t2, -t2
      w.J[(row0 + 1) * nv + i] = jn - mu * jt1;
      w.J[(row0 + 2) * nv + i] = jn + mu * jt2;
      w.J[(row0 + 3) * nv + i] = jn - mu * jt2;
    }
  }
  // 3. per-row impedance and regularisation
  for (int r = g.gl; r < nefc; r += G) {
    double K, B, imp, diag, Rr;
    const int info = w.row_info[r], rtype = row_type_of(info);
    if (rtype == ROW_LIMIT) {
      int j = row_id_of(info);
      make_kbip(m, &m.jnt_solref[2 * j], &m.jnt_solimp[5 * j], w.efc_pos[r], w.efc_margin[r], K, B, imp);
      diag = m.dof_invweight0[m.jnt_dofadr[j]];
      Rr = fmax(kMinVal, (1.0 - imp) * diag / imp);
    } else {
      int c = row_id_of(info), p = w.con_pair[c] & 0xffff;
      int g1 = m.pair_geom1[p], g2 = m.pair_geom2[p];
      ConParam cp;
      contact_param(m, g1, g2, cp);
      make_kbip(m, cp.solref, cp.solimp, w.efc_pos[r], w.efc_margin[r], K, B, imp);
      int b1 = m.geom_bodyid[g1], b2 = m.geom_bodyid[g2];
      double tran = m.body_invweight0[2 * b1] + m.body_invweight0[2 * b2];
      if (rtype == ROW_FRICTIONLESS) {
        Rr = fmax(kMinVal, (1.0 - imp) * tran / imp);
      } else {
        // all pyramid edges share R = 2 mu^2 R(first edge); the first edge has diagApprox = tran + mu^2 tran
        diag = tran + cp.mu * cp.mu * tran;
        double R0 = fmax(kMinVal, (1.0 - imp) * diag / imp);
        Rr = 2.0 * cp.mu * cp.mu * R0;
      }
    }
    w.efc_B[r] = B;
    w.efc_kterm[r] = K * imp * (w.efc_pos[r] - w.efc_margin[r]);
    w.efc_D[r] = 1.0 / Rr;
  }
  g.sync();
}

// mj_referenceConstraint: aref = -B (J qvel) - K imp (pos - margin)
template <int G, int NV>
TK_DEV void reference_constraint(const DM& m, WS& w, const Grp<G, NV>& g) {
  const int nv = NV ? NV : m.nv;
  for (int r = g.gl; r < w.nefc; r += G) {
    double vel = 0.0;
    for (int i = 0; i < nv; i++) vel += w.J[r * nv + i] * w.qvel[i];
    w.efc_aref[r] = -w.efc_B[r] * vel - w.efc_kterm[r];
  }
  g.sync();
}

}  // namespace tk
