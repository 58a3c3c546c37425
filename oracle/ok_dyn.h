// ORACLE (test infrastructure only).  Velocity / acceleration stages, Newton solver, Euler step:
// restates MuJoCo's mj_comVel, mj_passive, mj_rne, mj_fwdActuation, mj_fwdAcceleration,
// mj_fwdConstraint + mj_solNewton (engine_solver.c: PrimalSearch / PrimalEval / updateBracket),
// mj_EulerSkip, mj_forwardSkip and the fork's exported mj_stepSkip (engine_derivative_fd.c);
// reference call sites: src/Differentiator/Differentiator.cpp:69,118,155,253,276,357,377,
// src/Optimiser/iLQR.cpp:222,914.  PARITY UNPINNED (see ok_sim.h).
#pragma once
#include "ok_constraint.h"

namespace ok {

// mj_comVel
template <class R>
void com_vel(const Model* m, Data<R>* d) {
  zero(&d->cvel[0], 6);
  for (int i = 1; i < m->nbody; i++) {
    R cvel[6];
    copy(cvel, &d->cvel[6 * m->body_parentid[i]], 6);
    int bda = m->body_dofadr[i], dofnum = m->body_dofnum[i];
    int j = 0;
    while (j < dofnum) {
      int dj = bda + j;
      int jt = m->jnt_type[m->dof_jntid[dj]];
      if (jt == JNT_FREE) {
        zero(&d->cdof_dot[6 * dj], 18);
        for (int k = 0; k < 3; k++)
          for (int c = 0; c < 6; c++) cvel[c] += d->cdof[6 * (dj + k) + c] * d->qvel[dj + k];
        j += 3;
        dj = bda + j;
        for (int k = 0; k < 3; k++) cross_motion(&d->cdof_dot[6 * (dj + k)], cvel, &d->cdof[6 * (dj + k)]);
        for (int k = 0; k < 3; k++)
          for (int c = 0; c < 6; c++) cvel[c] += d->cdof[6 * (dj + k) + c] * d->qvel[dj + k];
        j += 3;
      } else {
        cross_motion(&d->cdof_dot[6 * dj], cvel, &d->cdof[6 * dj]);
        for (int c = 0; c < 6; c++) cvel[c] += d->cdof[6 * dj + c] * d->qvel[dj];
        j += 1;
      }
    }
    copy(&d->cvel[6 * i], cvel, 6);
  }
}

// mj_passive: joint damping + gravity compensation (no springs / fluid in scope)
template <class R>
void passive(const Model* m, Data<R>* d) {
  int nv = m->nv;
  for (int i = 0; i < nv; i++) d->qfrc_passive[i] = -R(m->dof_damping[i]) * d->qvel[i];
  // gravcomp: force -gravity*mass*gravcomp applied at the body com
  std::vector<R> jac(3 * nv);
  for (int b = 1; b < m->nbody; b++) {
    if (m->body_gravcomp[b] == 0.0) continue;
    R f[3];
    for (int k = 0; k < 3; k++) f[k] = -R(m->gravity[k]) * R(m->body_mass[b]) * R(m->body_gravcomp[b]);
    jac_point(m, d, jac.data(), &d->xipos[3 * b], b);
    for (int i = 0; i < nv; i++) d->qfrc_passive[i] += jac[i] * f[0] + jac[nv + i] * f[1] + jac[2 * nv + i] * f[2];
  }
}

// mj_rne with flg_acc = 0: bias forces (Coriolis, centrifugal, gravity)
template <class R>
void rne_bias(const Model* m, Data<R>* d, R* result) {
  int nb = m->nbody, nv = m->nv;
  std::vector<R> cacc(6 * nb, R(0.0)), cfrc(6 * nb, R(0.0));
  for (int k = 0; k < 3; k++) cacc[3 + k] = -R(m->gravity[k]);
  for (int i = 1; i < nb; i++) {
    int bda = m->body_dofadr[i];
    R tmp[6];
    zero(tmp, 6);
    for (int j = 0; j < m->body_dofnum[i]; j++)
      for (int c = 0; c < 6; c++) tmp[c] += d->cdof_dot[6 * (bda + j) + c] * d->qvel[bda + j];
    for (int c = 0; c < 6; c++) cacc[6 * i + c] = cacc[6 * m->body_parentid[i] + c] + tmp[c];
    R t1[6], t2[6], t3[6];
    mul_inert_vec(t1, &d->cinert[10 * i], &cacc[6 * i]);
    mul_inert_vec(t2, &d->cinert[10 * i], &d->cvel[6 * i]);
    cross_force(t3, &d->cvel[6 * i], t2);
    for (int c = 0; c < 6; c++) cfrc[6 * i + c] = t1[c] + t3[c];
  }
  for (int i = nb - 1; i > 0; i--) {
    int p = m->body_parentid[i];
    if (p) for (int c = 0; c < 6; c++) cfrc[6 * p + c] += cfrc[6 * i + c];
  }
  for (int i = 0; i < nv; i++) result[i] = dotn(&d->cdof[6 * i], &cfrc[6 * m->dof_bodyid[i]], 6);
}

// mj_fwdActuation (joint transmissions, no activation dynamics)
template <class R>
void fwd_actuation(const Model* m, Data<R>* d) {
  zero(d->qfrc_actuator.data(), m->nv);
  for (int i = 0; i < m->nu; i++) {
    R ctrl = d->ctrl[i];
    if (m->actuator_ctrllimited[i]) {
      R lo(m->actuator_ctrlrange[2 * i]), hi(m->actuator_ctrlrange[2 * i + 1]);
      if (ctrl < lo) ctrl = lo;
      if (ctrl > hi) ctrl = hi;
    }
    int dof = m->actuator_dofadr[i];
    R gear(m->actuator_gear[i]);
    R force = R(m->actuator_gainprm[i]) * ctrl;
    if (m->actuator_biastype[i] == 1) {
      int qadr = m->jnt_qposadr[m->actuator_trnid[i]];
      R length = gear * d->qpos[qadr], velocity = gear * d->qvel[dof];
      force += R(m->actuator_biasprm[3 * i]) + R(m->actuator_biasprm[3 * i + 1]) * length + R(m->actuator_biasprm[3 * i + 2]) * velocity;
    }
    if (m->actuator_forcelimited[i]) {
      R lo(m->actuator_forcerange[2 * i]), hi(m->actuator_forcerange[2 * i + 1]);
      if (force < lo) force = lo;
      if (force > hi) force = hi;
    }
    d->qfrc_actuator[dof] += gear * force;
  }
}

// mj_fwdAcceleration
template <class R>
void fwd_acceleration(const Model* m, Data<R>* d) {
  int nv = m->nv;
  for (int i = 0; i < nv; i++) {
    d->qfrc_smooth[i] = d->qfrc_passive[i] - d->qfrc_bias[i] + d->qfrc_actuator[i];
    d->qacc_smooth[i] = d->qfrc_smooth[i];
  }
  solve_ldl(m, d->qLD.data(), d->qLDiagInv.data(), d->qacc_smooth.data());
}

// ------------------------------- Newton solver --------------------------------------------------

// mj_constraintUpdate for limit / frictionless / pyramidal rows: force, state, cost
template <class R>
R constraint_update(const Model* m, Data<R>* d, const R* jar, bool write) {
  int nv = m->nv;
  R cost(0.0);
  for (int i = 0; i < d->nefc; i++) {
    if (jar[i] < R(0.0)) {
      cost += R(0.5) * d->efc_D[i] * jar[i] * jar[i];
      if (write) { d->efc_force[i] = -d->efc_D[i] * jar[i]; d->efc_state[i] = 1; }
    } else if (write) {
      d->efc_force[i] = R(0.0);
      d->efc_state[i] = 0;
    }
  }
  if (write) {
    for (int j = 0; j < nv; j++) {
      R s(0.0);
      for (int i = 0; i < d->nefc; i++)
        if (d->efc_force[i] != R(0.0)) s += d->efc_J[i * nv + j] * d->efc_force[i];
      d->qfrc_constraint[j] = s;
    }
  }
  return cost;
}

template <class R>
struct PrimalPnt { R alpha, cost, deriv[2]; };

template <class R>
struct PrimalCtx {
  std::vector<R> Jaref, Jv, Ma, Mv, grad, Mgrad, search, quad, H, L;
  R gauss, cost, quadGauss[3];
  int LSiter;
};

// PrimalEval for the constraint types in scope
template <class R>
void primal_eval(const Data<R>* d, PrimalCtx<R>* ctx, PrimalPnt<R>* p, R alpha) {
  ctx->LSiter++;
  R qt[3] = {ctx->quadGauss[0], ctx->quadGauss[1], ctx->quadGauss[2]};
  for (int i = 0; i < d->nefc; i++) {
    R x = ctx->Jaref[i] + alpha * ctx->Jv[i];
    if (x < R(0.0)) { qt[0] += ctx->quad[3 * i]; qt[1] += ctx->quad[3 * i + 1]; qt[2] += ctx->quad[3 * i + 2]; }
  }
  p->alpha = alpha;
  p->cost = alpha * alpha * qt[2] + alpha * qt[1] + qt[0];
  p->deriv[0] = R(2.0) * alpha * qt[2] + qt[1];
  p->deriv[1] = R(2.0) * qt[2];
  if (p->deriv[1] <= R(0.0)) p->deriv[1] = R(kMinVal);
}

template <class R>
int update_bracket(const Data<R>* d, PrimalCtx<R>* ctx, PrimalPnt<R>* p, const PrimalPnt<R>* cand, PrimalPnt<R>* pnext) {
  int flag = 0;
  for (int i = 0; i < 3; i++) {
    if (p->deriv[0] < R(0.0) && cand[i].deriv[0] < R(0.0) && p->deriv[0] < cand[i].deriv[0]) { *p = cand[i]; flag = 1; }
    else if (p->deriv[0] > R(0.0) && cand[i].deriv[0] > R(0.0) && p->deriv[0] > cand[i].deriv[0]) { *p = cand[i]; flag = 2; }
  }
  if (flag) primal_eval(d, ctx, pnext, p->alpha - p->deriv[0] / p->deriv[1]);
  return flag;
}

// PrimalSearch: exact line search along ctx->search
template <class R>
R primal_search(const Model* m, const Data<R>* d, PrimalCtx<R>* ctx, double tolerance) {
  using std::sqrt;
  int nv = m->nv, nefc = d->nefc;
  ctx->LSiter = 0;
  R snorm = sqrt(dotn(ctx->search.data(), ctx->search.data(), nv));
  if (snorm < R(kMinVal)) return R(0.0);
  R scale = R(1.0) / R(m->meaninertia * std::max(1, nv));
  R gtol = R(tolerance * m->ls_tolerance) * snorm / scale;
  // Mv, Jv
  mul_M(m, d, ctx->Mv.data(), ctx->search.data());
  for (int i = 0; i < nefc; i++) ctx->Jv[i] = dotn(&d->efc_J[i * nv], ctx->search.data(), nv);
  // PrimalPrepare
  ctx->quadGauss[0] = ctx->gauss;
  ctx->quadGauss[1] = dotn(ctx->search.data(), ctx->Ma.data(), nv) - dotn(d->qfrc_smooth.data(), ctx->search.data(), nv);
  ctx->quadGauss[2] = R(0.5) * dotn(ctx->search.data(), ctx->Mv.data(), nv);
  for (int i = 0; i < nefc; i++) {
    R DJ0 = d->efc_D[i] * ctx->Jaref[i];
    ctx->quad[3 * i] = R(0.5) * ctx->Jaref[i] * DJ0;
    ctx->quad[3 * i + 1] = ctx->Jv[i] * DJ0;
    ctx->quad[3 * i + 2] = R(0.5) * ctx->Jv[i] * d->efc_D[i] * ctx->Jv[i];
  }
  PrimalPnt<R> p0, p1, p2, pmid, p1next, p2next;
  primal_eval(d, ctx, &p0, R(0.0));
  primal_eval(d, ctx, &p1, p0.alpha - p0.deriv[0] / p0.deriv[1]);
  if (p0.cost < p1.cost) p1 = p0;
  if (fabs(p1.deriv[0]) < gtol) return p1.alpha;
  int dir = (p1.deriv[0] < R(0.0)) ? +1 : -1;
  int p2update = 0;
  p2 = p1;
  while (p1.deriv[0] * R((double)dir) <= -gtol && ctx->LSiter < m->ls_iterations) {
    p2 = p1;
    p2update = 1;
    primal_eval(d, ctx, &p1, p1.alpha - p1.deriv[0] / p1.deriv[1]);
    if (fabs(p1.deriv[0]) < gtol) return p1.alpha;
  }
  if (ctx->LSiter >= m->ls_iterations) return p1.alpha;
  if (!p2update) return p1.alpha;
  p2next = p1;
  primal_eval(d, ctx, &p1next, p1.alpha - p1.deriv[0] / p1.deriv[1]);
  while (ctx->LSiter < m->ls_iterations) {
    primal_eval(d, ctx, &pmid, R(0.5) * (p1.alpha + p2.alpha));
    PrimalPnt<R> cand[3] = {p1next, p2next, pmid};
    R bestcost(0.0);
    int bestind = -1;
    for (int i = 0; i < 3; i++)
      if (fabs(cand[i].deriv[0]) < gtol && (bestind == -1 || cand[i].cost < bestcost)) { bestcost = cand[i].cost; bestind = i; }
    if (bestind >= 0) return cand[bestind].alpha;
    int b1 = update_bracket(d, ctx, &p1, cand, &p1next);
    int b2 = update_bracket(d, ctx, &p2, cand, &p2next);
    if (!b1 && !b2) return pmid.alpha;
  }
  if (p1.cost <= p2.cost && p1.cost < p0.cost) return p1.alpha;
  if (p2.cost <= p1.cost && p2.cost < p0.cost) return p2.alpha;
  return R(0.0);
}

// dense Cholesky H = L L' (lower, row-major); mju_cholFactor with mjMINVAL floor
template <class R>
void chol_factor(R* L, const R* H, int n) {
  using std::sqrt;
  copy(L, H, n * n);
  for (int j = 0; j < n; j++) {
    R tmp = L[j * n + j];
    for (int k = 0; k < j; k++) tmp -= L[j * n + k] * L[j * n + k];
    if (tmp < R(kMinVal)) tmp = R(kMinVal);
    L[j * n + j] = sqrt(tmp);
    R inv = R(1.0) / L[j * n + j];
    for (int i = j + 1; i < n; i++) {
      R t = L[i * n + j];
      for (int k = 0; k < j; k++) t -= L[i * n + k] * L[j * n + k];
      L[i * n + j] = t * inv;
    }
  }
}
template <class R>
void chol_solve(R* x, const R* L, const R* b, int n) {
  for (int i = 0; i < n; i++) {
    R t = b[i];
    for (int k = 0; k < i; k++) t -= L[i * n + k] * x[k];
    x[i] = t / L[i * n + i];
  }
  for (int i = n - 1; i >= 0; i--) {
    R t = x[i];
    for (int k = i + 1; k < n; k++) t -= L[k * n + i] * x[k];
    x[i] = t / L[i * n + i];
  }
}

// mj_solNewton (pyramidal / frictionless / limit rows): H = M + J' diag(D*active) J refactored every
// iteration (MuJoCo updates the factor incrementally; the factor is mathematically identical)
template <class R>
void sol_newton(const Model* m, Data<R>* d, const SolverOpts& opt) {
  using std::sqrt;
  int nv = m->nv, nefc = d->nefc;
  PrimalCtx<R> ctx;
  ctx.Jaref.assign(nefc, R(0.0)); ctx.Jv.assign(nefc, R(0.0)); ctx.quad.assign(3 * nefc, R(0.0));
  for (auto* v : {&ctx.Ma, &ctx.Mv, &ctx.grad, &ctx.Mgrad, &ctx.search}) v->assign(nv, R(0.0));
  ctx.H.assign(nv * nv, R(0.0)); ctx.L.assign(nv * nv, R(0.0));
  mul_M(m, d, ctx.Ma.data(), d->qacc.data());
  for (int i = 0; i < nefc; i++) ctx.Jaref[i] = dotn(&d->efc_J[i * nv], d->qacc.data(), nv) - d->efc_aref[i];

  auto update_constraint = [&]() {
    ctx.cost = constraint_update(m, d, ctx.Jaref.data(), true);
    R g(0.0);
    for (int i = 0; i < nv; i++) g += (ctx.Ma[i] - d->qfrc_smooth[i]) * (d->qacc[i] - d->qacc_smooth[i]);
    ctx.gauss = R(0.5) * g;
    ctx.cost += ctx.gauss;
  };
  auto update_gradient = [&]() {
    for (int i = 0; i < nv; i++) ctx.grad[i] = ctx.Ma[i] - d->qfrc_smooth[i] - d->qfrc_constraint[i];
    copy(ctx.H.data(), d->qM.data(), nv * nv);
    for (int r = 0; r < nefc; r++) {
      if (!d->efc_state[r]) continue;
      const R* J = &d->efc_J[r * nv];
      for (int i = 0; i < nv; i++) {
        if (J[i] == R(0.0)) continue;
        R DJi = d->efc_D[r] * J[i];
        for (int j = 0; j <= i; j++) ctx.H[i * nv + j] += DJi * J[j];
      }
    }
    for (int i = 0; i < nv; i++)
      for (int j = i + 1; j < nv; j++) ctx.H[i * nv + j] = ctx.H[j * nv + i];
    chol_factor(ctx.L.data(), ctx.H.data(), nv);
    chol_solve(ctx.Mgrad.data(), ctx.L.data(), ctx.grad.data(), nv);
  };

  update_constraint();
  update_gradient();
  for (int i = 0; i < nv; i++) ctx.search[i] = -ctx.Mgrad[i];
  R scale = R(1.0) / R(m->meaninertia * std::max(1, nv));
  int iter = 0;
  while (iter < opt.iterations) {
    R alpha = primal_search(m, d, &ctx, opt.tolerance);
    if (alpha == R(0.0)) break;
    for (int i = 0; i < nv; i++) { d->qacc[i] += alpha * ctx.search[i]; ctx.Ma[i] += alpha * ctx.Mv[i]; }
    for (int i = 0; i < nefc; i++) ctx.Jaref[i] += alpha * ctx.Jv[i];
    R oldcost = ctx.cost;
    update_constraint();
    update_gradient();
    for (int i = 0; i < nv; i++) ctx.search[i] = -ctx.Mgrad[i];
    R improvement = scale * (oldcost - ctx.cost);
    R gradient = scale * sqrt(dotn(ctx.grad.data(), ctx.grad.data(), nv));
    iter++;
    if (improvement < R(opt.tolerance) || gradient < R(opt.tolerance)) break;
  }
  d->solver_niter = iter;
}

// mj_fwdConstraint
template <class R>
void fwd_constraint(const Model* m, Data<R>* d, const SolverOpts& opt) {
  int nv = m->nv, nefc = d->nefc;
  if (nefc == 0) {
    copy(d->qacc.data(), d->qacc_smooth.data(), nv);
    copy(d->qacc_warmstart.data(), d->qacc_smooth.data(), nv);
    zero(d->qfrc_constraint.data(), nv);
    d->solver_niter = 0;
    return;
  }
  std::vector<R> jar(nefc), Ma(nv);
  for (int i = 0; i < nefc; i++) d->efc_b[i] = dotn(&d->efc_J[i * nv], d->qacc_smooth.data(), nv) - d->efc_aref[i];
  // warm start: keep qacc_warmstart only if it is cheaper than qacc_smooth
  copy(d->qacc.data(), d->qacc_warmstart.data(), nv);
  for (int i = 0; i < nefc; i++) jar[i] = dotn(&d->efc_J[i * nv], d->qacc.data(), nv) - d->efc_aref[i];
  R cost_warm = constraint_update(m, d, jar.data(), false);
  mul_M(m, d, Ma.data(), d->qacc.data());
  R g(0.0);
  for (int i = 0; i < nv; i++) g += (Ma[i] - d->qfrc_smooth[i]) * (d->qacc[i] - d->qacc_smooth[i]);
  cost_warm += R(0.5) * g;
  R cost_smooth = constraint_update(m, d, d->efc_b.data(), false);
  if (cost_warm > cost_smooth) copy(d->qacc.data(), d->qacc_smooth.data(), nv);
  sol_newton(m, d, opt);
  copy(d->qacc_warmstart.data(), d->qacc.data(), nv);
}

// ------------------------------- integration ----------------------------------------------------

// mj_integratePos
template <class R>
void integrate_pos(const Model* m, R* qpos, const R* qvel, R dt) {
  for (int j = 0; j < m->njnt; j++) {
    int pa = m->jnt_qposadr[j], va = m->jnt_dofadr[j];
    switch (m->jnt_type[j]) {
      case JNT_FREE:
        for (int k = 0; k < 3; k++) qpos[pa + k] += dt * qvel[va + k];
        quat_integrate(qpos + pa + 3, qvel + va + 3, dt);
        break;
      case JNT_HINGE:
      case JNT_SLIDE:
        qpos[pa] += dt * qvel[va];
        break;
      default: break;
    }
  }
}

// mj_differentiatePos: qvel = (qpos2 - qpos1) / dt on the configuration manifold
template <class R>
void differentiate_pos(const Model* m, R* qvel, R dt, const R* qpos1, const R* qpos2) {
  for (int j = 0; j < m->njnt; j++) {
    int pa = m->jnt_qposadr[j], va = m->jnt_dofadr[j];
    switch (m->jnt_type[j]) {
      case JNT_FREE: {
        for (int k = 0; k < 3; k++) qvel[va + k] = (qpos2[pa + k] - qpos1[pa + k]) / dt;
        R v[3];
        sub_quat(v, qpos1 + pa + 3, qpos2 + pa + 3);
        for (int k = 0; k < 3; k++) qvel[va + 3 + k] = v[k] / dt;
        break;
      }
      case JNT_HINGE:
      case JNT_SLIDE:
        qvel[va] = (qpos2[pa] - qpos1[pa]) / dt;
        break;
      default: break;
    }
  }
}

// mj_EulerSkip: semi-implicit Euler with implicit joint damping
template <class R>
void euler_skip(const Model* m, Data<R>* d, bool skipfactor) {
  int nv = m->nv;
  R h(m->timestep);
  bool damping = false;
  for (int i = 0; i < nv; i++)
    if (m->dof_damping[i] > 0) { damping = true; break; }
  std::vector<R> qacc(nv);
  if (!damping) {
    copy(qacc.data(), d->qacc.data(), nv);
  } else {
    if (!skipfactor) {
      copy(d->qH.data(), d->qM.data(), nv * nv);
      for (int i = 0; i < nv; i++) d->qH[i * nv + i] += h * R(m->dof_damping[i]);
      std::vector<R> tmp(d->qH);
      factor_ldl(m, tmp.data(), d->qH.data(), d->qHDiagInv.data());
    }
    for (int i = 0; i < nv; i++) qacc[i] = d->qfrc_smooth[i] + d->qfrc_constraint[i];
    solve_ldl(m, d->qH.data(), d->qHDiagInv.data(), qacc.data());
  }
  // mj_advance
  for (int i = 0; i < nv; i++) d->qvel[i] += qacc[i] * h;
  integrate_pos(m, d->qpos.data(), d->qvel.data(), h);
  d->time += h;
}

template <class R>
bool bad(R x) { using std::isfinite; return !isfinite(x) || x > R(1e10) || x < R(-1e10); }

// mj_forwardSkip (sensors never computed: skip_sensor = 1 at every reference call site)
template <class R>
void forward_skip(const Model* m, Data<R>* d, int skipstage, const SolverOpts& opt) {
  if (skipstage < STAGE_POS) {
    kinematics(m, d);
    com_pos(m, d);
    crb(m, d);
    factor_ldl(m, d->qM.data(), d->qLD.data(), d->qLDiagInv.data());
    collision(m, d);
    make_constraint(m, d);
  }
  if (skipstage < STAGE_VEL) {
    com_vel(m, d);
    passive(m, d);
    reference_constraint(m, d);
    rne_bias(m, d, d->qfrc_bias.data());
  }
  fwd_actuation(m, d);
  fwd_acceleration(m, d);
  fwd_constraint(m, d, opt);
}

// mj_stepSkip (mj_step == skipstage NONE).  MuJoCo's mj_check* reset the state on NaN / huge values
// and raise a warning; the oracle records the warning bit and carries on (documented difference).
template <class R>
void step_skip(const Model* m, Data<R>* d, int skipstage, const SolverOpts& opt) {
  for (int i = 0; i < m->nq; i++) if (bad(d->qpos[i])) d->warning |= 1;
  for (int i = 0; i < m->nv; i++) if (bad(d->qvel[i])) d->warning |= 2;
  forward_skip(m, d, skipstage, opt);
  for (int i = 0; i < m->nv; i++) if (bad(d->qacc[i])) d->warning |= 4;
  euler_skip(m, d, skipstage >= STAGE_POS);
}

}  // namespace ok
