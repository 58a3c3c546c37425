// Position stage of the batched step: kinematics, centre-of-mass quantities, composite inertia,
// mass matrix + L'DL factor, collision, constraint rows.  (MuJoCo: mj_fwdPosition.)
#pragma once
#include "step.cuh"

namespace tk {

// mj_kinematics: level-synchronous over the body tree (lanes = bodies of one depth)
template <int G, int NV>
TK_DEV void kinematics(const DM& m, WS& w, const Grp<G, NV>& g) {
  if (g.gl == 0) {
    w.xpos[0] = w.xpos[1] = w.xpos[2] = 0.0;
    w.xquat[0] = 1.0; w.xquat[1] = w.xquat[2] = w.xquat[3] = 0.0;
    w.xipos[0] = w.xipos[1] = w.xipos[2] = 0.0;
  }
  g.sync();
  for (int depth = 1; depth <= m.treedepth; depth++) {
    for (int i = 1 + g.gl; i < m.nbody; i += G) {
      if (m.body_depth[i] != depth) continue;
      int pid = m.body_parentid[i];
      double xpos[3], xquat[4];
      int jntadr = m.body_jntadr[i], jntnum = m.body_jntnum[i];
      if (jntnum == 1 && m.jnt_type[jntadr] == JNT_FREE) {
        int qa = m.jnt_qposadr[jntadr];
        for (int k = 0; k < 3; k++) xpos[k] = w.qpos[qa + k];
        for (int k = 0; k < 4; k++) xquat[k] = w.qpos[qa + 3 + k];
        normalize4(xquat);
        for (int k = 0; k < 3; k++) w.xanchor[3 * jntadr + k] = xpos[k];
        w.xaxis[3 * jntadr] = 0.0; w.xaxis[3 * jntadr + 1] = 0.0; w.xaxis[3 * jntadr + 2] = 1.0;
      } else {
        double bpos[3] = {m.body_pos[3 * i], m.body_pos[3 * i + 1], m.body_pos[3 * i + 2]};
        double bquat[4] = {m.body_quat[4 * i], m.body_quat[4 * i + 1], m.body_quat[4 * i + 2], m.body_quat[4 * i + 3]};
        { double pm[9]; quat2mat(pm, &w.xquat[4 * pid]); mat_vec3(xpos, pm, bpos); }
        for (int k = 0; k < 3; k++) xpos[k] += w.xpos[3 * pid + k];
        mul_quat(xquat, &w.xquat[4 * pid], bquat);
        for (int j = jntadr; j < jntadr + jntnum; j++) {
          int qa = m.jnt_qposadr[j];
          double jaxis[3] = {m.jnt_axis[3 * j], m.jnt_axis[3 * j + 1], m.jnt_axis[3 * j + 2]};
          double jpos[3] = {m.jnt_pos[3 * j], m.jnt_pos[3 * j + 1], m.jnt_pos[3 * j + 2]};
          double xanchor[3], xaxis[3];
          rot_vec_quat(xaxis, jaxis, xquat);
          rot_vec_quat(xanchor, jpos, xquat);
          for (int k = 0; k < 3; k++) xanchor[k] += xpos[k];
          int jt = m.jnt_type[j];
          if (jt == JNT_SLIDE) {
            double dq = w.qpos[qa] - m.qpos0[qa];
            for (int k = 0; k < 3; k++) xpos[k] += xaxis[k] * dq;
          } else if (jt == JNT_HINGE) {
            double qloc[4], tmp[4], vec[3];
            axis_angle2quat(qloc, jaxis, w.qpos[qa] - m.qpos0[qa]);
            mul_quat(tmp, xquat, qloc);
            for (int k = 0; k < 4; k++) xquat[k] = tmp[k];
            rot_vec_quat(vec, jpos, xquat);
            for (int k = 0; k < 3; k++) xpos[k] = xanchor[k] - vec[k];
          }
          for (int k = 0; k < 3; k++) { w.xanchor[3 * j + k] = xanchor[k]; w.xaxis[3 * j + k] = xaxis[k]; }
        }
      }
      normalize4(xquat);
      for (int k = 0; k < 3; k++) w.xpos[3 * i + k] = xpos[k];
      for (int k = 0; k < 4; k++) w.xquat[4 * i + k] = xquat[k];
      double xm[9];
      quat2mat(xm, xquat);
      double ipos[3] = {m.body_ipos[3 * i], m.body_ipos[3 * i + 1], m.body_ipos[3 * i + 2]};
      double v[3];
      mat_vec3(v, xm, ipos);
      for (int k = 0; k < 3; k++) w.xipos[3 * i + k] = xpos[k] + v[k];
    }
    g.sync();
    TK_CLK(w, g, 23 + (depth < 4 ? depth : 4));
  }
}

// world pose of a geom (mj_kinematics' geom_xpos / geom_xmat), rebuilt from the body pose on demand
TK_DEV void geom_pose(const DM& m, const WS& w, int gi, double* gpos_out, double* gmat_out) {
  int b = m.geom_bodyid[gi];
  double gpos[3] = {m.geom_pos[3 * gi], m.geom_pos[3 * gi + 1], m.geom_pos[3 * gi + 2]};
  double gquat[4] = {m.geom_quat[4 * gi], m.geom_quat[4 * gi + 1], m.geom_quat[4 * gi + 2], m.geom_quat[4 * gi + 3]};
  double v[3], q[4], xm[9];
  if (b == 0) { for (int k = 0; k < 9; k++) xm[k] = (k % 4 == 0) ? 1.0 : 0.0; }
  else quat2mat(xm, &w.xquat[4 * b]);
  mat_vec3(v, xm, gpos);
  for (int k = 0; k < 3; k++) gpos_out[k] = w.xpos[3 * b + k] + v[k];
  mul_quat(q, &w.xquat[4 * b], gquat);
  quat2mat(gmat_out, q);
}

// mj_comPos: centre of mass of every kinematic tree (stored at its root body), com-frame inertias, motion axes
template <int G, int NV>
TK_DEV void com_pos(const DM& m, WS& w, const Grp<G, NV>& g) {
  // subtree com of root bodies (children of the world): the only subtree_com entries the pipeline reads
  for (int b = 1 + g.gl; b < m.nbody; b += G) {
    if (m.body_parentid[b] != 0) continue;
    unsigned lo = m.body_subtreemask_lo[b], hi = m.body_subtreemask_hi[b];
    double c[3] = {0.0, 0.0, 0.0}, mass = 0.0;
    // accumulate from the last body to the first: the order of MuJoCo's backward pass for a chain
    for (int j = m.nbody - 1; j >= b; j--) {
      if (!in_mask(lo, hi, j)) continue;
      double mj = m.body_mass[j];
      c[0] += w.xipos[3 * j] * mj; c[1] += w.xipos[3 * j + 1] * mj; c[2] += w.xipos[3 * j + 2] * mj;
      mass += mj;
    }
    if (mass < kMinVal) { c[0] = w.xipos[3 * b]; c[1] = w.xipos[3 * b + 1]; c[2] = w.xipos[3 * b + 2]; }
    else { c[0] /= mass; c[1] /= mass; c[2] /= mass; }
    w.rootcom[3 * b] = c[0]; w.rootcom[3 * b + 1] = c[1]; w.rootcom[3 * b + 2] = c[2];
  }
  g.sync();
  for (int i = g.gl; i < m.nbody; i += G) {
    double* res = &w.cinert[10 * i];
    if (i == 0) { for (int k = 0; k < 10; k++) res[k] = 0.0; continue; }
    int root = m.body_rootid[i];
    double dif[3], mass = m.body_mass[i];
    for (int k = 0; k < 3; k++) dif[k] = w.xipos[3 * i + k] - w.rootcom[3 * root + k];
    double mat[9];
    {
      double iquat[4] = {m.body_iquat[4 * i], m.body_iquat[4 * i + 1], m.body_iquat[4 * i + 2], m.body_iquat[4 * i + 3]}, q[4];
      mul_quat(q, &w.xquat[4 * i], iquat);
      quat2mat(mat, q);
    }
    double in0 = m.body_inertia[3 * i], in1 = m.body_inertia[3 * i + 1], in2 = m.body_inertia[3 * i + 2];
    double t0 = mat[0] * in0, t3 = mat[1] * in1, t6 = mat[2] * in2;
    double t1 = mat[3] * in0, t4 = mat[4] * in1, t7 = mat[5] * in2;
    double t2 = mat[6] * in0, t5 = mat[7] * in1, t8 = mat[8] * in2;
    res[0] = mat[0] * t0 + mat[1] * t3 + mat[2] * t6 + mass * (dif[1] * dif[1] + dif[2] * dif[2]);
    res[1] = mat[3] * t1 + mat[4] * t4 + mat[5] * t7 + mass * (dif[0] * dif[0] + dif[2] * dif[2]);
    res[2] = mat[6] * t2 + mat[7] * t5 + mat[8] * t8 + mass * (dif[0] * dif[0] + dif[1] * dif[1]);
    res[3] = mat[0] * t1 + mat[1] * t4 + mat[2] * t7 - mass * dif[0] * dif[1];
    res[4] = mat[0] * t2 + mat[1] * t5 + mat[2] * t8 - mass * dif[0] * dif[2];
    res[5] = mat[3] * t2 + mat[4] * t5 + mat[5] * t8 - mass * dif[1] * dif[2];
    res[6] = mass * dif[0]; res[7] = mass * dif[1]; res[8] = mass * dif[2];
    res[9] = mass;
  }
  for (int j = g.gl; j < m.njnt; j += G) {
    int bi = m.jnt_bodyid[j], da = m.jnt_dofadr[j], root = m.body_rootid[bi];
    double off[3];
    for (int k = 0; k < 3; k++) off[k] = w.rootcom[3 * root + k] - w.xanchor[3 * j + k];
    int jt = m.jnt_type[j];
    if (jt == JNT_FREE) {
      for (int k = 0; k < 18; k++) w.cdof[6 * da + k] = 0.0;
      w.cdof[6 * da + 3] = 1.0; w.cdof[6 * (da + 1) + 4] = 1.0; w.cdof[6 * (da + 2) + 5] = 1.0;
      double xm[9];
      quat2mat(xm, &w.xquat[4 * bi]);
      for (int k = 0; k < 3; k++) {
        double ax[3] = {xm[k], xm[3 + k], xm[6 + k]};
        double* c = &w.cdof[6 * (da + 3 + k)];
        c[0] = ax[0]; c[1] = ax[1]; c[2] = ax[2];
        cross3(c + 3, ax, off);
      }
    } else if (jt == JNT_SLIDE) {
      double* c = &w.cdof[6 * da];
      c[0] = c[1] = c[2] = 0.0;
      c[3] = w.xaxis[3 * j]; c[4] = w.xaxis[3 * j + 1]; c[5] = w.xaxis[3 * j + 2];
    } else {
      double* c = &w.cdof[6 * da];
      const double* ax = &w.xaxis[3 * j];
      c[0] = ax[0]; c[1] = ax[1]; c[2] = ax[2];
      cross3(c + 3, ax, off);
    }
  }
  g.sync();
}

// mj_crb + dense M, then L'DL factor of M (mj_factorM)
template <int G, int NV>
TK_DEV void crb_factor(const DM& m, WS& w, const Grp<G, NV>& g) {
  const int nv = NV ? NV : m.nv;
  for (int b = g.gl; b < m.nbody; b += G) {
    double acc[10];
    for (int k = 0; k < 10; k++) acc[k] = w.cinert[10 * b + k];
    if (b > 0) {
      unsigned lo = m.body_subtreemask_lo[b], hi = m.body_subtreemask_hi[b];
      for (int j = b + 1; j < m.nbody; j++) {
        if (!in_mask(lo, hi, j)) continue;
        for (int k = 0; k < 10; k++) acc[k] += w.cinert[10 * j + k];
      }
    }
    for (int k = 0; k < 10; k++) w.crb[10 * b + k] = acc[k];
  }
  for (int e = g.gl; e < nv * nv; e += G) w.M[e] = 0.0;
  g.sync();
  for (int i = g.gl; i < nv; i += G) {
    double buf[6];
    mul_inert_vec(buf, &w.crb[10 * m.dof_bodyid[i]], &w.cdof[6 * i]);
    anc_t<NV> chain = anc_mask<NV>(m, i);
    while (chain) {
      const int j = pop_top(chain);
      const double* c = &w.cdof[6 * j];
      double v = c[0] * buf[0] + c[1] * buf[1] + c[2] * buf[2] + c[3] * buf[3] + c[4] * buf[4] + c[5] * buf[5];
      if (j == i) w.M[i * nv + i] = m.dof_armature[i] + v;
      else { w.M[i * nv + j] = v; w.M[j * nv + i] = v; }
    }
  }
  g.sync();
}

// L'DL factorisation (elimination from the last dof upwards, ancestors only): LD <- factor(A)
template <int G, int NV>
TK_DEV void factor_ldl(const DM& m, const double* A, double* LD, double* dinv, const Grp<G, NV>& g) {
  const int nv = NV ? NV : m.nv;
  for (int e = g.gl; e < nv * nv; e += G) LD[e] = A[e];
  g.sync();
  if (NV == 0 && m.max_tree_dofs < nv) {
    // several kinematic trees (an arm and free bodies): M is block diagonal over them and the elimination of one
    // tree never touches another, so trip t eliminates dof treelast - t of EVERY tree at once - as many trips as the
    // largest tree has dofs instead of nv, every entry through the same operations as in the sequential order
#pragma unroll 1
    for (int t = 0; t < m.max_tree_dofs; t++) {
      double a0 = 0.0;
      bool have0 = false;
      for (int i = g.gl; i < nv; i += G) {
        const int k = m.dof_treelast[i] - t;
        if (k <= i) continue;
        if (!((anc_mask<NV>(m, k) >> i) & 1)) continue;
        double Mkk = LD[k * nv + k];
        if (Mkk < kMinVal) Mkk = kMinVal;
        double a = LD[k * nv + i] / Mkk;
        if (i == g.gl) { a0 = a; have0 = true; }
        anc_t<NV> chain = anc_mask<NV>(m, i);
        while (chain) {
          const int j = pop_top(chain);
          LD[i * nv + j] -= a * LD[k * nv + j];
        }
      }
      g.sync();
      for (int i = g.gl; i < nv; i += G) {
        const int k = m.dof_treelast[i] - t;
        if (k == i) {  // this trip's dof of the tree: clamp the pivot, keep its inverse
          double Mkk = LD[k * nv + k];
          if (Mkk < kMinVal) Mkk = kMinVal;
          LD[k * nv + k] = Mkk; dinv[k] = 1.0 / Mkk;
        } else if (k > i && ((anc_mask<NV>(m, k) >> i) & 1)) {
          if (i == g.gl && have0) {
            LD[k * nv + i] = a0;
          } else {
            double Mkk = LD[k * nv + k];
            if (Mkk < kMinVal) Mkk = kMinVal;
            LD[k * nv + i] = LD[k * nv + i] / Mkk;
          }
        }
      }
      g.sync();
    }
    return;
  }
#pragma unroll 1
  for (int k = nv - 1; k >= 0; k--) {
    double Mkk = LD[k * nv + k];
    if (Mkk < kMinVal) Mkk = kMinVal;
    const anc_t<NV> anck = anc_mask<NV>(m, k);
    // each ancestor i of k updates its own row; row k is only read.  The quotient is also the new LD[k][i]: the lane
    // keeps it for its first row and writes it once the other lanes have read the old value
    double a0 = 0.0;
    bool have0 = false;
    for (int i = g.gl; i < k; i += G) {
      if (!((anck >> i) & 1)) continue;
      double a = LD[k * nv + i] / Mkk;
      if (i == g.gl) { a0 = a; have0 = true; }
      anc_t<NV> chain = anc_mask<NV>(m, i);
      while (chain) {
        const int j = pop_top(chain);
        LD[i * nv + j] -= a * LD[k * nv + j];
      }
    }
    g.sync();
    if (have0) LD[k * nv + g.gl] = a0;
    for (int i = g.gl + G; i < k; i += G)
      if ((anck >> i) & 1) LD[k * nv + i] = LD[k * nv + i] / Mkk;
    if (g.gl == 0) { LD[k * nv + k] = Mkk; dinv[k] = 1.0 / Mkk; }
    g.sync();
  }
}

// x <- inv(L'DL) x   (x in shared memory)
template <int G, int NV>
TK_DEV void solve_ldl(const DM& m, const double* LD, const double* dinv, double* x, const Grp<G, NV>& g) {
  const int nv = NV ? NV : m.nv;
  // x <- inv(L') x : once x[i] is final, push it to its ancestors (several trees: one dof of every tree per trip, see
  // factor_ldl)
  if (NV == 0 && m.max_tree_dofs < nv) {
#pragma unroll 1
    for (int t = 0; t < m.max_tree_dofs - 1; t++) {
      for (int j = g.gl; j < nv; j += G) {
        const int i = m.dof_treelast[j] - t;
        if (i > j && ((anc_mask<NV>(m, i) >> j) & 1)) x[j] -= LD[i * nv + j] * x[i];
      }
      g.sync();
    }
  } else {
#pragma unroll 1
    for (int i = nv - 1; i > 0; i--) {
      double xi = x[i];
      if (xi != 0.0) {
        const anc_t<NV> anci = anc_mask<NV>(m, i);
        for (int j = g.gl; j < i; j += G)
          if ((anci >> j) & 1) x[j] -= LD[i * nv + j] * xi;
      }
      g.sync();
    }
  }
  for (int i = g.gl; i < nv; i += G) x[i] *= dinv[i];
  g.sync();
  // x <- inv(L) x : row i needs the final values of its ancestors.  Level-synchronous over the depth of the dof in the
  // tree (= its number of ancestors): all dofs of one depth are independent, each lane walks its own parent chain in
  // the order of the sequential algorithm, so the values are the same bits.
  {
    int depth[(64 + G - 1) / G];  // nv <= 64 (ancestor masks have 64 bits)
    int maxd = 0;
#pragma unroll
    for (int k = 0; k < (64 + G - 1) / G; k++) {
      const int i = g.gl + k * G;
      depth[k] = i < nv ? __popc(m.dof_ancmask_lo[i]) + __popc(m.dof_ancmask_hi[i]) : -1;
      maxd = depth[k] > maxd ? depth[k] : maxd;
    }
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) {
      const int o = __shfl_xor_sync(g.mask, maxd, off, G);
      maxd = o > maxd ? o : maxd;
    }
#pragma unroll 1
    for (int d = 1; d <= maxd; d++) {
#pragma unroll
      for (int k = 0; k < (64 + G - 1) / G; k++) {
        if (depth[k] != d) continue;
        const int i = g.gl + k * G;
        double s = x[i];
        anc_t<NV> chain = anc_mask<NV>(m, i) & ~(anc_t<NV>(1) << i);
        while (chain) {
          const int j = pop_top(chain);
          s -= LD[i * nv + j] * x[j];
        }
        x[i] = s;
      }
      g.sync();
    }
  }
}

// y = M x
template <int G, int NV>
TK_DEV void mul_M(const DM& m, const WS& w, double* y, const double* x, const Grp<G, NV>& g) {
  const int nv = NV ? NV : m.nv;
  for (int i = g.gl; i < nv; i += G) {
    double s = 0.0;
    for (int j = 0; j < nv; j++) s += w.M[i * nv + j] * x[j];
    y[i] = s;
  }
}

// translational Jacobian column of a world point rigidly attached to `body`, for dof i (zero if i does not move body)
TK_DEV void jac_col(const DM& m, const WS& w, double* col, const double* point, int body, int i) {
  col[0] = col[1] = col[2] = 0.0;
  while (body > 0 && m.body_dofnum[body] == 0) body = m.body_parentid[body];
  if (body == 0) return;
  int last = m.body_dofadr[body] + m.body_dofnum[body] - 1;
  if (i > last || !in_mask(m.dof_ancmask_lo[last], m.dof_ancmask_hi[last], i)) return;
  int root = m.body_rootid[body];
  double off[3] = {point[0] - w.rootcom[3 * root], point[1] - w.rootcom[3 * root + 1], point[2] - w.rootcom[3 * root + 2]};
  const double* c = &w.cdof[6 * i];
  double t[3];
  cross3(t, c, off);
  col[0] = c[3] + t[0]; col[1] = c[4] + t[1]; col[2] = c[5] + t[2];
}

}  // namespace tk
