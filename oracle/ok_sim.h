// ORACLE (test infrastructure only - never on the product path).
// CPU restatement of the MuJoCo forward-dynamics step the reference calls
// (mj_step at src/Differentiator/Differentiator.cpp:69, src/Optimiser/iLQR.cpp:222,914 and the
//  fork's mj_stepSkip at Differentiator.cpp:118,155,253,276,357,377).  MuJoCo is a third-party
// dependency absent from /root/reference (custom fork DMackRus/mujoco, version unpinned, API level
// >= 3.0 - SURVEY.md F1-F3); what follows restates its published algorithm (MuJoCo "Computation"
// chapter + 3.1-series engine_*.c structure): kinematics, comPos, CRB, L'DL factor, collision,
// constraint construction (limits + pyramidal contacts), comVel, passive, RNE bias, actuation,
// smooth acceleration, Newton primal solver with exact line search, semi-implicit Euler.
// PARITY UNPINNED against real MuJoCo (no MuJoCo, no golden vectors in the reference).
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>

#include "ok_math.h"
#include "ok_model.h"

namespace ok {

enum { STAGE_NONE = 0, STAGE_POS = 1, STAGE_VEL = 2 };
enum { EFC_LIMIT = 0, EFC_CONTACT_FRICTIONLESS = 1, EFC_CONTACT_PYRAMIDAL = 2 };

template <class R>
struct Contact {
  R dist, pos[3], frame[9];
  R includemargin, friction[5], solref[2], solimp[5], mu;
  int dim, geom1, geom2, exclude, efc_address;
};

struct SolverOpts {
  int iterations;
  double tolerance;
};

template <class R>
struct Data {
  const Model* m;
  // state (the 8-field snapshot of MuJoCoHelper::CpMjData; applied forces are always zero in scope)
  R time;
  std::vector<R> qpos, qvel, qacc, qacc_warmstart, ctrl;
  // position stage
  std::vector<R> xpos, xquat, xmat, xipos, ximat, xanchor, xaxis, geom_xpos, geom_xmat;
  std::vector<R> subtree_com, cinert, cdof, crb, qM, qLD, qLDiagInv;
  std::vector<Contact<R>> contact;
  int nefc = 0, ncon = 0;
  std::vector<int> efc_type, efc_id;
  std::vector<R> efc_J, efc_pos, efc_margin, efc_diagApprox, efc_R, efc_D, efc_KBIP;
  // velocity stage
  std::vector<R> cvel, cdof_dot, qfrc_passive, qfrc_bias, efc_vel, efc_aref;
  // acceleration stage
  std::vector<R> qfrc_actuator, qfrc_smooth, qacc_smooth, efc_b, efc_force, qfrc_constraint;
  std::vector<int> efc_state;
  // Euler
  std::vector<R> qH, qHDiagInv;
  // diagnostics
  int solver_niter = 0;
  int warning = 0;  // bit0 bad qpos, bit1 bad qvel, bit2 bad qacc, bit3 contact/efc overflow

  explicit Data(const Model* model) : m(model) {
    int nq = m->nq, nv = m->nv, nb = m->nbody, nj = m->njnt, ng = m->ngeom, ne = m->maxefc;
    time = R(0.0);
    qpos.assign(nq, R(0.0)); qvel.assign(nv, R(0.0)); qacc.assign(nv, R(0.0)); qacc_warmstart.assign(nv, R(0.0));
    ctrl.assign(m->nu, R(0.0));
    for (int i = 0; i < nq; i++) qpos[i] = R(m->qpos0[i]);
    xpos.assign(3 * nb, R(0.0)); xquat.assign(4 * nb, R(0.0)); xmat.assign(9 * nb, R(0.0));
    xipos.assign(3 * nb, R(0.0)); ximat.assign(9 * nb, R(0.0));
    xanchor.assign(3 * nj, R(0.0)); xaxis.assign(3 * nj, R(0.0));
    geom_xpos.assign(3 * ng, R(0.0)); geom_xmat.assign(9 * ng, R(0.0));
    subtree_com.assign(3 * nb, R(0.0)); cinert.assign(10 * nb, R(0.0)); crb.assign(10 * nb, R(0.0));
    cdof.assign(6 * nv, R(0.0)); cdof_dot.assign(6 * nv, R(0.0));
    qM.assign(nv * nv, R(0.0)); qLD.assign(nv * nv, R(0.0)); qLDiagInv.assign(nv, R(0.0));
    qH.assign(nv * nv, R(0.0)); qHDiagInv.assign(nv, R(0.0));
    contact.resize(m->maxcontact);
    efc_type.assign(ne, 0); efc_id.assign(ne, 0); efc_state.assign(ne, 0);
    efc_J.assign(ne * nv, R(0.0));
    for (auto* v : {&efc_pos, &efc_margin, &efc_diagApprox, &efc_R, &efc_D, &efc_vel, &efc_aref, &efc_b, &efc_force})
      v->assign(ne, R(0.0));
    efc_KBIP.assign(4 * ne, R(0.0));
    cvel.assign(6 * nb, R(0.0));
    for (auto* v : {&qfrc_passive, &qfrc_bias, &qfrc_actuator, &qfrc_smooth, &qacc_smooth, &qfrc_constraint})
      v->assign(nv, R(0.0));
  }
};

// ============================== position stage ==================================================

// mj_kinematics
template <class R>
void kinematics(const Model* m, Data<R>* d) {
  d->xquat[0] = R(1.0);
  d->xmat[0] = R(1.0); d->xmat[4] = R(1.0); d->xmat[8] = R(1.0);
  d->ximat[0] = R(1.0); d->ximat[4] = R(1.0); d->ximat[8] = R(1.0);
  for (int i = 1; i < m->nbody; i++) {
    int pid = m->body_parentid[i];
    R xpos[3], xquat[4];
    R bpos[3] = {R(m->body_pos[3 * i]), R(m->body_pos[3 * i + 1]), R(m->body_pos[3 * i + 2])};
    R bquat[4] = {R(m->body_quat[4 * i]), R(m->body_quat[4 * i + 1]), R(m->body_quat[4 * i + 2]), R(m->body_quat[4 * i + 3])};
    int jntadr = m->body_jntadr[i], jntnum = m->body_jntnum[i];
    if (jntnum == 1 && m->jnt_type[jntadr] == JNT_FREE) {
      int qadr = m->jnt_qposadr[jntadr];
      for (int k = 0; k < 3; k++) xpos[k] = d->qpos[qadr + k];
      for (int k = 0; k < 4; k++) xquat[k] = d->qpos[qadr + 3 + k];
      normalize4(xquat);
      for (int k = 0; k < 3; k++) { d->xanchor[3 * jntadr + k] = xpos[k]; }
      d->xaxis[3 * jntadr] = R(0.0); d->xaxis[3 * jntadr + 1] = R(0.0); d->xaxis[3 * jntadr + 2] = R(1.0);
    } else {
      // position and orientation in the parent frame
      mul_mat_vec3(xpos, &d->xmat[9 * pid], bpos);
      for (int k = 0; k < 3; k++) xpos[k] += d->xpos[3 * pid + k];
      mul_quat(xquat, &d->xquat[4 * pid], bquat);
      for (int j = jntadr; j < jntadr + jntnum; j++) {
        int qadr = m->jnt_qposadr[j];
        R jaxis[3] = {R(m->jnt_axis[3 * j]), R(m->jnt_axis[3 * j + 1]), R(m->jnt_axis[3 * j + 2])};
        R jpos[3] = {R(m->jnt_pos[3 * j]), R(m->jnt_pos[3 * j + 1]), R(m->jnt_pos[3 * j + 2])};
        R xanchor[3], xaxis[3];
        rot_vec_quat(xaxis, jaxis, xquat);
        rot_vec_quat(xanchor, jpos, xquat);
        for (int k = 0; k < 3; k++) xanchor[k] += xpos[k];
        switch (m->jnt_type[j]) {
          case JNT_SLIDE: {
            R dq = d->qpos[qadr] - R(m->qpos0[qadr]);
            for (int k = 0; k < 3; k++) xpos[k] += xaxis[k] * dq;
            break;
          }
          case JNT_HINGE: {
            R qloc[4], tmp[4], vec[3];
            axis_angle2quat(qloc, jaxis, d->qpos[qadr] - R(m->qpos0[qadr]));
            mul_quat(tmp, xquat, qloc);
            for (int k = 0; k < 4; k++) xquat[k] = tmp[k];
            // correct for off-centre rotation
            rot_vec_quat(vec, jpos, xquat);
            for (int k = 0; k < 3; k++) xpos[k] = xanchor[k] - vec[k];
            break;
          }
          default:
            d->warning |= 16;
        }
        for (int k = 0; k < 3; k++) { d->xanchor[3 * j + k] = xanchor[k]; d->xaxis[3 * j + k] = xaxis[k]; }
      }
    }
    normalize4(xquat);
    for (int k = 0; k < 3; k++) d->xpos[3 * i + k] = xpos[k];
    for (int k = 0; k < 4; k++) d->xquat[4 * i + k] = xquat[k];
    quat2mat(&d->xmat[9 * i], xquat);
    // inertial frame
    R ipos[3] = {R(m->body_ipos[3 * i]), R(m->body_ipos[3 * i + 1]), R(m->body_ipos[3 * i + 2])};
    R iquat[4] = {R(m->body_iquat[4 * i]), R(m->body_iquat[4 * i + 1]), R(m->body_iquat[4 * i + 2]), R(m->body_iquat[4 * i + 3])};
    R v[3], q[4];
    mul_mat_vec3(v, &d->xmat[9 * i], ipos);
    for (int k = 0; k < 3; k++) d->xipos[3 * i + k] = xpos[k] + v[k];
    mul_quat(q, xquat, iquat);
    quat2mat(&d->ximat[9 * i], q);
  }
  for (int g = 0; g < m->ngeom; g++) {
    int b = m->geom_bodyid[g];
    R gpos[3] = {R(m->geom_pos[3 * g]), R(m->geom_pos[3 * g + 1]), R(m->geom_pos[3 * g + 2])};
    R gquat[4] = {R(m->geom_quat[4 * g]), R(m->geom_quat[4 * g + 1]), R(m->geom_quat[4 * g + 2]), R(m->geom_quat[4 * g + 3])};
    R v[3], q[4];
    mul_mat_vec3(v, &d->xmat[9 * b], gpos);
    for (int k = 0; k < 3; k++) d->geom_xpos[3 * g + k] = d->xpos[3 * b + k] + v[k];
    mul_quat(q, &d->xquat[4 * b], gquat);
    quat2mat(&d->geom_xmat[9 * g], q);
  }
}

// mj_comPos: subtree centres of mass, com-based body inertias and joint motion axes
template <class R>
void com_pos(const Model* m, Data<R>* d) {
  int nb = m->nbody;
  std::vector<R> mass_subtree(nb, R(0.0));
  for (int i = 0; i < nb; i++) {
    R mass(m->body_mass[i]);
    for (int k = 0; k < 3; k++) d->subtree_com[3 * i + k] = d->xipos[3 * i + k] * mass;
    mass_subtree[i] = mass;
  }
  for (int i = nb - 1; i > 0; i--) {
    int p = m->body_parentid[i];
    for (int k = 0; k < 3; k++) d->subtree_com[3 * p + k] += d->subtree_com[3 * i + k];
    mass_subtree[p] += mass_subtree[i];
  }
  for (int i = 0; i < nb; i++) {
    if (mass_subtree[i] < R(kMinVal)) {
      for (int k = 0; k < 3; k++) d->subtree_com[3 * i + k] = d->xipos[3 * i + k];
    } else {
      for (int k = 0; k < 3; k++) d->subtree_com[3 * i + k] /= mass_subtree[i];
    }
  }
  zero(&d->cinert[0], 10);
  for (int i = 1; i < nb; i++) {
    int root = m->body_rootid[i];
    R offset[3], inert[3] = {R(m->body_inertia[3 * i]), R(m->body_inertia[3 * i + 1]), R(m->body_inertia[3 * i + 2])};
    for (int k = 0; k < 3; k++) offset[k] = d->xipos[3 * i + k] - d->subtree_com[3 * root + k];
    inert_com(&d->cinert[10 * i], inert, &d->ximat[9 * i], offset, R(m->body_mass[i]));
  }
  for (int j = 0; j < m->njnt; j++) {
    int bi = m->jnt_bodyid[j], da = m->jnt_dofadr[j], root = m->body_rootid[bi];
    R offset[3];
    for (int k = 0; k < 3; k++) offset[k] = d->subtree_com[3 * root + k] - d->xanchor[3 * j + k];
    switch (m->jnt_type[j]) {
      case JNT_FREE: {
        zero(&d->cdof[6 * da], 18);
        d->cdof[6 * da + 3] = R(1.0);
        d->cdof[6 * (da + 1) + 4] = R(1.0);
        d->cdof[6 * (da + 2) + 5] = R(1.0);
        for (int k = 0; k < 3; k++) {
          R axis[3] = {d->xmat[9 * bi + k], d->xmat[9 * bi + 3 + k], d->xmat[9 * bi + 6 + k]};
          dof_com(&d->cdof[6 * (da + 3 + k)], axis, offset);
        }
        break;
      }
      case JNT_SLIDE:
        dof_com<R>(&d->cdof[6 * da], &d->xaxis[3 * j], nullptr);
        // slide: linear motion along the axis
        for (int k = 0; k < 3; k++) { d->cdof[6 * da + 3 + k] = d->xaxis[3 * j + k]; d->cdof[6 * da + k] = R(0.0); }
        break;
      case JNT_HINGE:
        dof_com(&d->cdof[6 * da], &d->xaxis[3 * j], offset);
        break;
      default: break;
    }
  }
}

// mj_crb: composite rigid body inertia -> joint-space inertia (dense, both triangles filled)
template <class R>
void crb(const Model* m, Data<R>* d) {
  int nv = m->nv, nb = m->nbody;
  copy(&d->crb[0], &d->cinert[0], 10 * nb);
  for (int i = nb - 1; i > 0; i--) {
    int p = m->body_parentid[i];
    if (p > 0) for (int k = 0; k < 10; k++) d->crb[10 * p + k] += d->crb[10 * i + k];
  }
  zero(&d->qM[0], nv * nv);
  for (int i = 0; i < nv; i++) {
    R buf[6];
    mul_inert_vec(buf, &d->crb[10 * m->dof_bodyid[i]], &d->cdof[6 * i]);
    d->qM[i * nv + i] = R(m->dof_armature[i]);
    for (int j = i; j >= 0; j = m->dof_parentid[j]) {
      R v = dotn(&d->cdof[6 * j], buf, 6);
      if (j == i) d->qM[i * nv + i] += v;
      else { d->qM[i * nv + j] = v; d->qM[j * nv + i] = v; }
    }
  }
}

// mj_factorI: L'*D*L factorisation exploiting only the order of elimination MuJoCo uses
// (from the last dof upwards).  `LD` holds L below the diagonal (row k, columns < k) and D on it.
template <class R>
void factor_ldl(const Model* m, const R* M, R* LD, R* diaginv) {
  int nv = m->nv;
  copy(LD, M, nv * nv);
  for (int k = nv - 1; k >= 0; k--) {
    R Mkk = LD[k * nv + k];
    if (Mkk < R(kMinVal)) { Mkk = R(kMinVal); LD[k * nv + k] = Mkk; }
    R inv = R(1.0) / Mkk;
    diaginv[k] = inv;
    for (int i = m->dof_parentid[k]; i >= 0; i = m->dof_parentid[i]) {
      R a = LD[k * nv + i] / Mkk;
      for (int j = i; j >= 0; j = m->dof_parentid[j]) LD[i * nv + j] -= a * LD[k * nv + j];
      LD[k * nv + i] = a;
    }
  }
}

// mj_solveLD: x <- inv(L'*D*L) x
template <class R>
void solve_ldl(const Model* m, const R* LD, const R* diaginv, R* x) {
  int nv = m->nv;
  for (int i = nv - 1; i >= 0; i--) {
    if (x[i] == R(0.0)) continue;
    for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j]) x[j] -= LD[i * nv + j] * x[i];
  }
  for (int i = 0; i < nv; i++) x[i] *= diaginv[i];
  for (int i = 0; i < nv; i++)
    for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j]) x[i] -= LD[i * nv + j] * x[j];
}

// y = M x (dense symmetric storage)
template <class R>
void mul_M(const Model* m, const Data<R>* d, R* y, const R* x) {
  int nv = m->nv;
  for (int i = 0; i < nv; i++) {
    R s(0.0);
    // MuJoCo's sparse mj_mulM touches only ancestor/descendant pairs; the others are exact zeros
    for (int j = 0; j < nv; j++)
      if (d->qM[i * nv + j] != R(0.0)) s += d->qM[i * nv + j] * x[j];
    y[i] = s;
  }
}

// mj_jac restricted to the translational Jacobian of a world point attached to `body`:
// jacp[3 x nv] (row-major)
template <class R>
void jac_point(const Model* m, const Data<R>* d, R* jacp, const R* point, int body) {
  int nv = m->nv;
  zero(jacp, 3 * nv);
  // skip fixed bodies
  while (body > 0 && m->body_dofnum[body] == 0) body = m->body_parentid[body];
  if (body == 0) return;
  int root = m->body_rootid[body];
  R offset[3];
  for (int k = 0; k < 3; k++) offset[k] = point[k] - d->subtree_com[3 * root + k];
  int i = m->body_dofadr[body] + m->body_dofnum[body] - 1;
  while (i >= 0) {
    const R* c = &d->cdof[6 * i];
    R tmp[3];
    cross(tmp, c, offset);
    for (int k = 0; k < 3; k++) jacp[k * nv + i] = c[3 + k] + tmp[k];
    i = m->dof_parentid[i];
  }
}

}  // namespace ok
