// ORACLE (test infrastructure only).  Collision detection and constraint construction:
// restates MuJoCo's mj_collision (engine_collision_driver.c / engine_collision_primitive.c:
// mjraw_PlaneSphere, mjc_PlaneCapsule, mjc_PlaneBox) and mj_makeConstraint
// (engine_core_constraint.c: mj_instantiateLimit, mj_instantiateContact, mj_diagApprox,
// mj_makeImpedance, mj_referenceConstraint) for joint limits and pyramidal/frictionless contacts.
// MuJoCo is not in /root/reference; call sites on the reference path: every mj_step / mj_stepSkip
// (src/Differentiator/Differentiator.cpp:69,118,...; src/Optimiser/iLQR.cpp:222,914).
#pragma once
#include "ok_sim.h"

namespace ok {

constexpr double kMinImp = 0.0001, kMaxImp = 0.9999;

// mjraw_PlaneSphere
template <class R>
int plane_sphere(Contact<R>* con, R margin, const R* pos1, const R* mat1, const R* pos2, R radius) {
  R normal[3] = {mat1[2], mat1[5], mat1[8]};
  R tmp[3] = {pos2[0] - pos1[0], pos2[1] - pos1[1], pos2[2] - pos1[2]};
  R cdist = dot3(tmp, normal);
  if (cdist > margin + radius) return 0;
  con->dist = cdist - radius;
  R s = -con->dist * R(0.5) - radius;
  for (int k = 0; k < 3; k++) { con->pos[k] = pos2[k] + normal[k] * s; con->frame[k] = normal[k]; }
  for (int k = 3; k < 9; k++) con->frame[k] = R(0.0);
  return 1;
}

// mjc_PlaneCapsule
template <class R>
int plane_capsule(Contact<R>* con, R margin, const R* pos1, const R* mat1, const R* pos2, const R* mat2, const double* size2) {
  R axis[3] = {mat2[2], mat2[5], mat2[8]}, seg[3], p[3];
  for (int k = 0; k < 3; k++) seg[k] = axis[k] * R(size2[1]);
  for (int k = 0; k < 3; k++) p[k] = pos2[k] + seg[k];
  int n1 = plane_sphere(con, margin, pos1, mat1, p, R(size2[0]));
  for (int k = 0; k < 3; k++) p[k] = pos2[k] - seg[k];
  int n2 = plane_sphere(con + n1, margin, pos1, mat1, p, R(size2[0]));
  // align contact frames with the capsule axis
  if (n1) for (int k = 0; k < 3; k++) con[0].frame[3 + k] = axis[k];
  if (n2) for (int k = 0; k < 3; k++) con[n1].frame[3 + k] = axis[k];
  return n1 + n2;
}

// mjc_PlaneBox: corners below margin, at most 4
template <class R>
int plane_box(Contact<R>* con, R margin, const R* pos1, const R* mat1, const R* pos2, const R* mat2, const double* size2) {
  R norm[3] = {mat1[2], mat1[5], mat1[8]};
  R dif[3] = {pos2[0] - pos1[0], pos2[1] - pos1[1], pos2[2] - pos1[2]};
  R dist = dot3(dif, norm);
  int cnt = 0;
  for (int i = 0; i < 8; i++) {
    R vec[3], corner[3];
    vec[0] = R((i & 1) ? size2[0] : -size2[0]);
    vec[1] = R((i & 2) ? size2[1] : -size2[1]);
    vec[2] = R((i & 4) ? size2[2] : -size2[2]);
    mul_mat_vec3(corner, mat2, vec);
    R ldist = dot3(norm, corner);
    if (dist + ldist > margin || ldist > R(0.0)) continue;
    R cdist = dist + ldist;
    con[cnt].dist = cdist;
    for (int k = 0; k < 3; k++) {
      con[cnt].pos[k] = corner[k] + pos2[k] - norm[k] * cdist * R(0.5);
      con[cnt].frame[k] = norm[k];
    }
    for (int k = 3; k < 9; k++) con[cnt].frame[k] = R(0.0);
    if (++cnt >= 4) return 4;
  }
  return cnt;
}

// contact parameter mixing (mj_contactParam)
template <class R>
void contact_param(const Model* m, Contact<R>* con, int g1, int g2, R* margin, R* gap) {
  *margin = R(std::max(m->geom_margin[g1], m->geom_margin[g2]));
  *gap = R(std::max(m->geom_gap[g1], m->geom_gap[g2]));
  int p1 = m->geom_priority[g1], p2 = m->geom_priority[g2];
  double fri[3];
  if (p1 != p2) {
    int g = p1 > p2 ? g1 : g2;
    con->dim = m->geom_condim[g];
    for (int k = 0; k < 2; k++) con->solref[k] = R(m->geom_solref[2 * g + k]);
    for (int k = 0; k < 5; k++) con->solimp[k] = R(m->geom_solimp[5 * g + k]);
    for (int k = 0; k < 3; k++) fri[k] = m->geom_friction[3 * g + k];
  } else {
    con->dim = std::max(m->geom_condim[g1], m->geom_condim[g2]);
    double s1 = m->geom_solmix[g1], s2 = m->geom_solmix[g2], mix;
    if (s1 >= kMinVal && s2 >= kMinVal) mix = s1 / (s1 + s2);
    else if (s1 < kMinVal && s2 < kMinVal) mix = 0.5;
    else if (s1 < kMinVal) mix = 0.0;
    else mix = 1.0;
    const double* r1 = &m->geom_solref[2 * g1];
    const double* r2 = &m->geom_solref[2 * g2];
    if (r1[0] > 0 && r2[0] > 0) {
      for (int k = 0; k < 2; k++) con->solref[k] = R(mix * r1[k] + (1 - mix) * r2[k]);
    } else {
      for (int k = 0; k < 2; k++) con->solref[k] = R(std::min(r1[k], r2[k]));
    }
    for (int k = 0; k < 5; k++) con->solimp[k] = R(mix * m->geom_solimp[5 * g1 + k] + (1 - mix) * m->geom_solimp[5 * g2 + k]);
    for (int k = 0; k < 3; k++) fri[k] = std::max(m->geom_friction[3 * g1 + k], m->geom_friction[3 * g2 + k]);
  }
  con->friction[0] = R(fri[0]); con->friction[1] = R(fri[0]); con->friction[2] = R(fri[1]);
  con->friction[3] = R(fri[2]); con->friction[4] = R(fri[2]);
}

template <class R>
int plane_convex(const Model* m, Contact<R>* con, R margin, const R* pos1, const R* mat1, int g2, const R* pos2, const R* mat2);
template <class R>
int convex_convex(const Model* m, Contact<R>* con, R margin, int g1, const R* pos1, const R* mat1, int g2, const R* pos2,
                  const R* mat2);
template <class R>
int plane_cylinder(Contact<R>* con, R margin, const R* pos1, const R* mat1, const R* pos2, const R* mat2, const double* size2);
inline bool is_convex_type(int t) {
  return t == GEOM_SPHERE || t == GEOM_CAPSULE || t == GEOM_CYLINDER || t == GEOM_BOX || t == GEOM_MESH;
}

// mj_collision over the statically filtered pair list
template <class R>
void collision(const Model* m, Data<R>* d) {
  d->ncon = 0;
  for (int p = 0; p < m->npair; p++) {
    int g1 = m->pair_geom1[p], g2 = m->pair_geom2[p];
    int t1 = m->geom_type[g1], t2 = m->geom_type[g2];
    Contact<R> tmp[8];
    R margin, gap;
    contact_param(m, &tmp[0], g1, g2, &margin, &gap);
    // bounding-sphere test (plane: distance of the sphere centre to the plane)
    const R* pos1 = &d->geom_xpos[3 * g1];
    const R* mat1 = &d->geom_xmat[9 * g1];
    const R* pos2 = &d->geom_xpos[3 * g2];
    const R* mat2 = &d->geom_xmat[9 * g2];
    int n = 0;
    if (t1 == GEOM_PLANE) {
      R normal[3] = {mat1[2], mat1[5], mat1[8]};
      R dif[3] = {pos2[0] - pos1[0], pos2[1] - pos1[1], pos2[2] - pos1[2]};
      if (dot3(dif, normal) > margin + R(m->geom_rbound[g2])) continue;
      if (t2 == GEOM_SPHERE) n = plane_sphere(tmp, margin, pos1, mat1, pos2, R(m->geom_size[3 * g2]));
      else if (t2 == GEOM_CAPSULE) n = plane_capsule(tmp, margin, pos1, mat1, pos2, mat2, &m->geom_size[3 * g2]);
      else if (t2 == GEOM_BOX) n = plane_box(tmp, margin, pos1, mat1, pos2, mat2, &m->geom_size[3 * g2]);
      else if (t2 == GEOM_CYLINDER) n = plane_cylinder(tmp, margin, pos1, mat1, pos2, mat2, &m->geom_size[3 * g2]);
      else if (t2 == GEOM_MESH) n = plane_convex(m, tmp, margin, pos1, mat1, g2, pos2, mat2);
      else { d->warning |= 32; continue; }
    } else if (is_convex_type(t1) && is_convex_type(t2) &&
               (t1 == GEOM_MESH || t2 == GEOM_MESH || (t1 == GEOM_CYLINDER && t2 == GEOM_CYLINDER) ||
                (t1 == GEOM_CYLINDER && t2 == GEOM_BOX) || (t1 == GEOM_BOX && t2 == GEOM_CYLINDER))) {
      // mj_collideGeoms: bounding-sphere filter, then the general convex collider (mjc_Convex)
      R dif[3] = {pos2[0] - pos1[0], pos2[1] - pos1[1], pos2[2] - pos1[2]};
      R bound = R(m->geom_rbound[g1]) + R(m->geom_rbound[g2]) + margin;
      if (dot3(dif, dif) > bound * bound) continue;
      n = convex_convex(m, tmp, margin, g1, pos1, mat1, g2, pos2, mat2);
    } else {
      d->warning |= 32;  // pair type outside the implemented subset
      continue;
    }
    for (int i = 0; i < n; i++) {
      if (d->ncon >= m->maxcontact) { d->warning |= 8; break; }
      Contact<R>& c = d->contact[d->ncon];
      c = tmp[i];
      c.dim = tmp[0].dim;
      for (int k = 0; k < 5; k++) { c.friction[k] = tmp[0].friction[k]; c.solimp[k] = tmp[0].solimp[k]; }
      for (int k = 0; k < 2; k++) c.solref[k] = tmp[0].solref[k];
      c.geom1 = g1; c.geom2 = g2;
      c.includemargin = margin - gap;
      c.exclude = (c.dist >= c.includemargin) ? 1 : 0;
      make_frame(c.frame);
      c.efc_address = -1;
      d->ncon++;
    }
  }
}

// getimpedance + K,B from solref (mj_makeImpedance)
template <class R>
void make_kbip(const Model* m, const R* solref_in, const R* solimp_in, R pos, R margin, R* KBIP) {
  using std::pow;
  R solimp[5] = {solimp_in[0], solimp_in[1], solimp_in[2], solimp_in[3], solimp_in[4]};
  // clamp solimp (mj_assignImp)
  if (solimp[0] < R(kMinImp)) solimp[0] = R(kMinImp); if (solimp[0] > R(kMaxImp)) solimp[0] = R(kMaxImp);
  if (solimp[1] < R(kMinImp)) solimp[1] = R(kMinImp); if (solimp[1] > R(kMaxImp)) solimp[1] = R(kMaxImp);
  if (solimp[2] < R(0.0)) solimp[2] = R(0.0);
  if (solimp[3] < R(kMinImp)) solimp[3] = R(kMinImp); if (solimp[3] > R(kMaxImp)) solimp[3] = R(kMaxImp);
  if (solimp[4] < R(1.0)) solimp[4] = R(1.0);
  R solref[2] = {solref_in[0], solref_in[1]};
  // refsafe: time constant at least 2*timestep
  if (solref[0] > R(0.0) && solref[0] < R(2.0 * m->timestep)) solref[0] = R(2.0 * m->timestep);
  R dmin = solimp[0], dmax = solimp[1], width = solimp[2], mid = solimp[3], power = solimp[4];
  R imp, impP;
  R x = (pos - margin);
  if (!(dmin != dmax) || width < R(kMinVal)) {
    imp = R(0.5) * (dmin + dmax);
    impP = R(0.0);
  } else {
    x = x / width;
    R sgn(1.0);
    if (x < R(0.0)) { x = -x; sgn = R(-1.0); }
    if (x >= R(1.0) || x <= R(0.0)) {
      imp = (x >= R(1.0)) ? dmax : dmin;
      impP = R(0.0);
    } else {
      R y, yP;
      if (power == R(1.0)) { y = x; yP = R(1.0); }
      else if (x <= mid) {
        R a = R(1.0) / pow(mid, power - R(1.0));
        y = a * pow(x, power);
        yP = power * a * pow(x, power - R(1.0));
      } else {
        R b = R(1.0) / pow(R(1.0) - mid, power - R(1.0));
        y = R(1.0) - b * pow(R(1.0) - x, power);
        yP = power * b * pow(R(1.0) - x, power - R(1.0));
      }
      imp = dmin + y * (dmax - dmin);
      impP = yP * sgn * (dmax - dmin) / width;
    }
  }
  R K, B;
  if (solref[0] > R(0.0)) {
    K = R(1.0) / (dmax * dmax * solref[0] * solref[0] * solref[1] * solref[1]);
    B = R(2.0) / (dmax * solref[0]);
  } else {
    K = -solref[0] / (dmax * dmax);
    B = -solref[1] / dmax;
  }
  KBIP[0] = K; KBIP[1] = B; KBIP[2] = imp; KBIP[3] = impP;
}

// mj_makeConstraint: limits then contacts (no equality / friction-loss rows in scope)
template <class R>
void make_constraint(const Model* m, Data<R>* d) {
  int nv = m->nv;
  d->nefc = 0;
  auto add_row = [&](int type, int id, R pos, R margin) -> R* {
    if (d->nefc >= m->maxefc) { d->warning |= 8; return nullptr; }
    int r = d->nefc++;
    d->efc_type[r] = type; d->efc_id[r] = id; d->efc_pos[r] = pos; d->efc_margin[r] = margin;
    R* J = &d->efc_J[r * nv];
    zero(J, nv);
    return J;
  };
  // --- joint limits (mj_instantiateLimit)
  for (int j = 0; j < m->njnt; j++) {
    if (!m->jnt_limited[j]) continue;
    if (m->jnt_type[j] != JNT_HINGE && m->jnt_type[j] != JNT_SLIDE) continue;
    R value = d->qpos[m->jnt_qposadr[j]], margin(m->jnt_margin[j]);
    for (int side = -1; side <= 1; side += 2) {
      R dist = R((double)side) * (R(m->jnt_range[2 * j + (side + 1) / 2]) - value);
      if (dist < margin) {
        R* J = add_row(EFC_LIMIT, j, dist, margin);
        if (J) J[m->jnt_dofadr[j]] = R((double)-side);
      }
    }
  }
  // --- contacts (mj_instantiateContact), pyramidal cones
  std::vector<R> jac1(3 * nv), jac2(3 * nv), jdif(3 * nv);
  for (int c = 0; c < d->ncon; c++) {
    Contact<R>& con = d->contact[c];
    if (con.exclude) continue;
    int b1 = m->geom_bodyid[con.geom1], b2 = m->geom_bodyid[con.geom2];
    jac_point(m, d, jac1.data(), con.pos, b1);
    jac_point(m, d, jac2.data(), con.pos, b2);
    for (int k = 0; k < 3 * nv; k++) jdif[k] = jac2[k] - jac1[k];
    // rotate into the contact frame: row r = frame[r] . jdif
    std::vector<R> jc(3 * nv);
    for (int r = 0; r < 3; r++)
      for (int i = 0; i < nv; i++)
        jc[r * nv + i] = con.frame[3 * r] * jdif[i] + con.frame[3 * r + 1] * jdif[nv + i] + con.frame[3 * r + 2] * jdif[2 * nv + i];
    if (con.dim == 1) {
      con.efc_address = d->nefc;
      R* J = add_row(EFC_CONTACT_FRICTIONLESS, c, con.dist, con.includemargin);
      if (J) for (int i = 0; i < nv; i++) J[i] = jc[i];
    } else if (con.dim == 3) {
      con.efc_address = d->nefc;
      for (int k = 1; k < con.dim; k++) {
        for (int sgn = 1; sgn >= -1; sgn -= 2) {
          R* J = add_row(EFC_CONTACT_PYRAMIDAL, c, con.dist, con.includemargin);
          if (J)
            for (int i = 0; i < nv; i++) J[i] = jc[i] + R((double)sgn) * con.friction[k - 1] * jc[k * nv + i];
        }
      }
    } else {
      d->warning |= 32;  // condim 4/6 need rotational Jacobians: outside the implemented subset
    }
  }
  // --- diagApprox (mj_diagApprox)
  for (int i = 0; i < d->nefc; i++) {
    int id = d->efc_id[i];
    if (d->efc_type[i] == EFC_LIMIT) {
      d->efc_diagApprox[i] = R(m->dof_invweight0[m->jnt_dofadr[id]]);
    } else {
      const Contact<R>& con = d->contact[id];
      int b1 = m->geom_bodyid[con.geom1], b2 = m->geom_bodyid[con.geom2];
      R tran = R(m->body_invweight0[2 * b1]) + R(m->body_invweight0[2 * b2]);
      R rot = R(m->body_invweight0[2 * b1 + 1]) + R(m->body_invweight0[2 * b2 + 1]);
      if (d->efc_type[i] == EFC_CONTACT_FRICTIONLESS) d->efc_diagApprox[i] = tran;
      else {
        int j = i - con.efc_address;
        R fri = con.friction[j / 2];
        d->efc_diagApprox[i] = tran + fri * fri * (j < 4 ? tran : rot);
      }
    }
  }
  // --- impedance, R, D (mj_makeImpedance)
  for (int i = 0; i < d->nefc; i++) {
    int id = d->efc_id[i];
    R solref[2], solimp[5];
    if (d->efc_type[i] == EFC_LIMIT) {
      for (int k = 0; k < 2; k++) solref[k] = R(m->jnt_solref[2 * id + k]);
      for (int k = 0; k < 5; k++) solimp[k] = R(m->jnt_solimp[5 * id + k]);
    } else {
      for (int k = 0; k < 2; k++) solref[k] = d->contact[id].solref[k];
      for (int k = 0; k < 5; k++) solimp[k] = d->contact[id].solimp[k];
    }
    make_kbip(m, solref, solimp, d->efc_pos[i], d->efc_margin[i], &d->efc_KBIP[4 * i]);
    R imp = d->efc_KBIP[4 * i + 2];
    R r = (R(1.0) - imp) * d->efc_diagApprox[i] / imp;
    d->efc_R[i] = r < R(kMinVal) ? R(kMinVal) : r;
  }
  // pyramidal friction: common R for all basis vectors of a contact
  for (int c = 0; c < d->ncon; c++) {
    Contact<R>& con = d->contact[c];
    if (con.exclude || con.efc_address < 0 || con.dim == 1) continue;
    int i = con.efc_address;
    con.mu = con.friction[0];
    R Rpy = R(2.0) * con.mu * con.mu * d->efc_R[i];
    for (int j = 0; j < 2 * (con.dim - 1); j++) d->efc_R[i + j] = Rpy;
  }
  for (int i = 0; i < d->nefc; i++) d->efc_D[i] = R(1.0) / d->efc_R[i];
}

// mj_referenceConstraint: efc_vel = J qvel ; aref = -B vel - K imp (pos - margin)
template <class R>
void reference_constraint(const Model* m, Data<R>* d) {
  int nv = m->nv;
  for (int i = 0; i < d->nefc; i++) {
    d->efc_vel[i] = dotn(&d->efc_J[i * nv], d->qvel.data(), nv);
    const R* k = &d->efc_KBIP[4 * i];
    d->efc_aref[i] = -k[1] * d->efc_vel[i] - k[0] * k[2] * (d->efc_pos[i] - d->efc_margin[i]);
  }
}

}  // namespace ok

#include "ok_convex.h"
