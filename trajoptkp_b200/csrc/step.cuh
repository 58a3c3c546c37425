// Batched fp64 rigid-body step for sm_100a: one *group* of G lanes (G = 8, 16 or 32; G = 32 is
// warp-per-world) advances one world.  All mjData-equivalent state of the world lives in the
// group's shared-memory workspace; model arrays are read-only global (L1/L2 resident).
//
// Replaces MuJoCo's mj_step / mj_stepSkip on the reference path
// (src/Differentiator/Differentiator.cpp:69,118,155,253,276,357,377; src/Optimiser/iLQR.cpp:222,914).
// Pipeline order follows mj_forwardSkip: kinematics, comPos, crb, factorM, collision, makeConstraint |
// comVel, passive, referenceConstraint, rne | actuation, acceleration, Newton constraint solver | Euler.
// Lane mapping: tree stages are level-synchronous over bodies, everything else is strided over
// dofs / constraint rows / matrix entries; reductions are xor-butterflies inside the group, so every
// lane of a group holds bit-identical scalars and control flow is uniform per group.
#pragma once
#include <type_traits>
#include <cuda_runtime.h>

#include "device_model.h"

namespace tk {

#define TK_DEV __device__ __forceinline__
constexpr double kMinVal = 1e-15;
constexpr double kPi = 3.14159265358979323846;
constexpr double kMinImp = 0.0001, kMaxImp = 0.9999;
enum { JNT_FREE = 0, JNT_BALL = 1, JNT_SLIDE = 2, JNT_HINGE = 3 };
enum { GEOM_PLANE = 0, GEOM_SPHERE = 2, GEOM_CAPSULE = 3, GEOM_CYLINDER = 5, GEOM_BOX = 6, GEOM_MESH = 7 };
enum { ROW_LIMIT = 0, ROW_FRICTIONLESS = 1, ROW_PYRAMIDAL = 2 };
// status bits (per world)
enum { ST_BAD_QPOS = 1, ST_BAD_QVEL = 2, ST_BAD_QACC = 4, ST_OVERFLOW = 8, ST_UNSUPPORTED = 32 };

// stage clocks of one tracked world (development builds with -DTOKP_STAGE_CLOCKS only; tools/stage_clocks.py)
#ifdef TOKP_STAGE_CLOCKS
static __device__ unsigned long long g_stage_clk[32];
#define TK_CLK(w, g, i) do { if ((w).clk_on && (g).gl == 0) { long long t_ = clock64(); \
  atomicAdd(&tk::g_stage_clk[i], (unsigned long long)(t_ - (w).clk)); (w).clk = t_; } } while (0)
#define TK_CNT(w, g, i) do { if ((w).clk_on && (g).gl == 0) atomicAdd(&tk::g_stage_clk[i], 1ull); } while (0)
#else
#define TK_CLK(w, g, i) do {} while (0)
#define TK_CNT(w, g, i) do {} while (0)
#endif

// ---- group of G lanes -----------------------------------------------------------------------------
template <int G, int NV = 0>
struct Grp {
  // NV > 0: the model's nv is this compile-time constant (loops unroll, index arithmetic folds); 0 = read m.nv
  static constexpr int kNV = NV;
  unsigned mask;
  unsigned wmask;  // lanes of every group of this warp that runs a world (set by the kernel with a full-warp ballot)
  int gl;  // lane index inside the group
  // block-synchronous mode (set by the kernel, uniform over the block): 0 = warps run independently; >= 1 = every
  // thread of the block runs a world and the block passes the stages of a step together (one hot region of the large
  // step code in the instruction cache at a time); >= 2 also the Newton iterations; >= 3 also the line-search trips
  int bsync = 0;
  TK_DEV Grp() {
    int lane = threadIdx.x & 31;
    gl = lane % G;
    mask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (lane - gl));
    wmask = mask;
  }
  // warp-uniform loop condition: true while any running group of the warp still wants to iterate.  All running
  // groups must reach every any_warp() call; it doubles as the point where the groups of a warp reconverge, so that
  // the worlds sharing a warp execute the solver's loops in lock-step instead of one after the other.
  TK_DEV int any_warp(int p) const { return __any_sync(wmask, p); }
  // the same for a loop the whole block walks together in block-synchronous mode
  TK_DEV int any_block(int p) const { return bsync >= 2 ? __syncthreads_or(p) : __any_sync(wmask, p); }
  TK_DEV int any_block_ls(int p) const { return bsync >= 3 ? __syncthreads_or(p) : __any_sync(wmask, p); }
  // stage boundary of a step
  TK_DEV void phase() const { if (bsync) __syncthreads(); }
  TK_DEV void sync() const { __syncwarp(mask); }
  TK_DEV double sum(double v) const {
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1) v += __shfl_xor_sync(mask, v, off, G);
    return v;
  }
  // bit l = predicate of lane l of this group
  TK_DEV unsigned ballot(int p) const { return __ballot_sync(mask, p) >> ((threadIdx.x & 31) - gl); }
  TK_DEV int any(int p) const { return __any_sync(mask, p); }
  TK_DEV int all(int p) const { return __all_sync(mask, p); }
};

// ---- small vector helpers ---------------------------------------------------------------------------
TK_DEV double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
TK_DEV void cross3(double* r, const double* a, const double* b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
TK_DEV double normalize3(double* a) {
  double n = sqrt(dot3(a, a));
  if (n < kMinVal) { a[0] = 1.0; a[1] = 0.0; a[2] = 0.0; }
  else { double inv = 1.0 / n; a[0] *= inv; a[1] *= inv; a[2] *= inv; }
  return n;
}
TK_DEV void normalize4(double* q) {
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < kMinVal) { q[0] = 1.0; q[1] = 0.0; q[2] = 0.0; q[3] = 0.0; }
  else if (fabs(n - 1.0) > kMinVal) { double inv = 1.0 / n; q[0] *= inv; q[1] *= inv; q[2] *= inv; q[3] *= inv; }
}
TK_DEV void mul_quat(double* r, const double* a, const double* b) {
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
TK_DEV void quat2mat(double* m, const double* q) {
  double q00 = q[0] * q[0], q01 = q[0] * q[1], q02 = q[0] * q[2], q03 = q[0] * q[3];
  double q11 = q[1] * q[1], q12 = q[1] * q[2], q13 = q[1] * q[3];
  double q22 = q[2] * q[2], q23 = q[2] * q[3], q33 = q[3] * q[3];
  m[0] = q00 + q11 - q22 - q33; m[4] = q00 - q11 + q22 - q33; m[8] = q00 - q11 - q22 + q33;
  m[1] = 2.0 * (q12 - q03); m[2] = 2.0 * (q13 + q02);
  m[3] = 2.0 * (q12 + q03); m[5] = 2.0 * (q23 - q01);
  m[6] = 2.0 * (q13 - q02); m[7] = 2.0 * (q23 + q01);
}
TK_DEV void mat_vec3(double* r, const double* m, const double* v) {
  double x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  double y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  double z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
TK_DEV void rot_vec_quat(double* r, const double* v, const double* q) {
  if (q[0] == 1.0 && q[1] == 0.0 && q[2] == 0.0 && q[3] == 0.0) { r[0] = v[0]; r[1] = v[1]; r[2] = v[2]; return; }
  double m[9];
  quat2mat(m, q);
  mat_vec3(r, m, v);
}
TK_DEV void axis_angle2quat(double* q, const double* axis, double angle) {
  if (angle == 0.0) { q[0] = 1.0; q[1] = 0.0; q[2] = 0.0; q[3] = 0.0; return; }
  double s, c;
  sincos(angle * 0.5, &s, &c);
  q[0] = c; q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}
TK_DEV void quat_integrate(double* q, const double* vel, double scale) {
  double tmp[3] = {vel[0], vel[1], vel[2]};
  double angle = scale * normalize3(tmp);
  double qrot[4], res[4];
  axis_angle2quat(qrot, tmp, angle);
  normalize4(q);
  mul_quat(res, q, qrot);
  q[0] = res[0]; q[1] = res[1]; q[2] = res[2]; q[3] = res[3];
}
TK_DEV void sub_quat(double* res, const double* qa, const double* qb) {
  double qneg[4] = {qa[0], -qa[1], -qa[2], -qa[3]}, qdif[4];
  mul_quat(qdif, qneg, qb);
  double axis[3] = {qdif[1], qdif[2], qdif[3]};
  double s = normalize3(axis);
  double speed = 2.0 * atan2(s, qdif[0]);
  if (speed > kPi) speed -= 2.0 * kPi;
  res[0] = axis[0] * speed; res[1] = axis[1] * speed; res[2] = axis[2] * speed;
}
// spatial inertia (10 numbers) times motion vector [rot; lin]
TK_DEV void mul_inert_vec(double* res, const double* i, const double* v) {
  res[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
  res[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
  res[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
  res[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
  res[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
  res[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
}
TK_DEV void cross_motion(double* res, const double* vel, const double* v) {
  double a[3], b[3], c[3];
  cross3(a, vel, v);
  cross3(b, vel, v + 3);
  cross3(c, vel + 3, v);
  res[0] = a[0]; res[1] = a[1]; res[2] = a[2];
  res[3] = b[0] + c[0]; res[4] = b[1] + c[1]; res[5] = b[2] + c[2];
}
TK_DEV void cross_force(double* res, const double* vel, const double* f) {
  double a[3], b[3], c[3];
  cross3(a, vel, f);
  cross3(b, vel + 3, f + 3);
  cross3(c, vel, f + 3);
  res[0] = a[0] + b[0]; res[1] = a[1] + b[1]; res[2] = a[2] + b[2];
  res[3] = c[0]; res[4] = c[1]; res[5] = c[2];
}
TK_DEV bool in_mask(unsigned lo, unsigned hi, int bit) { return bit < 32 ? ((lo >> bit) & 1u) : ((hi >> (bit - 32)) & 1u); }
// The parent chain of a dof as a bit set: dof_ancmask has bit j set when j is the dof itself or one of its ancestors, and
// parents have smaller indices, so taking the set bits from the top visits i, parent(i), parent(parent(i)), ... in the
// order of a walk through dof_parentid - without a dependent global load per hop.
// The nv-specialised kernels with nv <= 32 keep the mask in one 32-bit register (clz and shifts are single instructions).
template <int NV>
using anc_t = typename std::conditional<(NV > 0 && NV <= 32), unsigned, unsigned long long>::type;
template <int NV>
TK_DEV anc_t<NV> anc_mask(const DM& m, int i) {
  if constexpr (NV > 0 && NV <= 32) return m.dof_ancmask_lo[i];
  else return ((unsigned long long)m.dof_ancmask_hi[i] << 32) | (unsigned long long)m.dof_ancmask_lo[i];
}
TK_DEV int pop_top(unsigned long long& mk) {
  int j = 63 - __clzll((long long)mk);
  mk &= ~(1ull << j);
  return j;
}
TK_DEV int pop_top(unsigned& mk) {
  int j = 31 - __clz((int)mk);
  mk &= ~(1u << j);
  return j;
}
TK_DEV bool bad_value(double x) { return !(fabs(x) <= 1e10); }

// ---- per-world workspace (shared memory) ---------------------------------------------------------
// Layout = persistent part P, followed by a union region shared by two groups of arrays with disjoint lifetimes:
//   A  position-stage temporaries (dead once make_constraint has produced J, D, B, kterm)
//      xpos xquat | (xanchor xaxis)  U  crb  U  (con_dist con_pos con_frame) | int tables
//   B  (cvel cacc cfrc cdof_dot : velocity stage)  U  (H jar jv efc_force [quad] : solver + Euler)
// Rotation matrices (xmat, ximat, geom_xmat) are not stored: they are rebuilt from xquat where needed with the same
// operation sequence (quat2mat then mat_vec3), so results are bit-identical to the stored-matrix form.
// `ecap` is the constraint-row capacity of this launch: the fast pass uses a small cap, worlds that overflow it are
// flagged (ST_OVERFLOW) and re-run by a second pass with ecap = maxefc.
struct WS {
  // state
  double *qpos, *qvel, *qacc, *warm, *ctrl;
  // position stage
  double *xpos, *xquat, *xipos, *xanchor, *xaxis;
  double *rootcom, *cinert, *crb, *cdof, *cdof_dot;
  double *M, *LD, *LDinv, *H, *Hinv;  // Hinv: diag inverse for the Euler factor
  // contacts
  double *con_dist, *con_pos, *con_frame;
  int *con_pair;
  // constraint rows
  double *J, *efc_pos, *efc_margin, *efc_D, *efc_aref, *efc_B, *efc_kterm, *jar, *jv, *quad, *efc_force;
  int *row_info, *cnt;  // row_info: packed (type | (sub+1) << 2 | id << 5); cnt: scratch counters (max(njnt, npair) + 2)
  // velocity / acceleration stage
  double *cvel, *cacc, *cfrc;
  double *frc_passive, *frc_bias, *frc_act, *frc_smooth, *acc_smooth, *frc_con, *Ma, *Mv, *grad, *Mgrad, *search, *tmpv;
  int ncon, nefc, ecap;
  int work;  // cost evaluations of the line searches of this world so far (a measure of its solver work)
#ifdef TOKP_STAGE_CLOCKS
  long long clk; int clk_on;
#endif
};

__host__ __device__ inline int ws_imax(int a, int b) { return a > b ? a : b; }

// rows of the line search that one lane keeps in registers (step_ls.cuh); with ecap <= LS_ROWS * lanes the quad[]
// array of the shared-memory path is never touched and gets no space
__host__ __device__ constexpr int ls_rows(int lanes) { return lanes >= 32 ? 1 : (lanes >= 16 ? 2 : 4); }
__host__ __device__ inline int ws_quad_doubles(int ecap, int lanes) { return ecap > ls_rows(lanes) * lanes ? 3 * ecap : 0; }

__host__ __device__ inline int ws_doubles(const DM& m, int ecap, int lanes) {
  int nb = m.nbody, nv = m.nv, ne = ecap, nc = m.maxcontact;
  int P = m.nq + 3 * nv + m.nu + 3 * nb + 3 * nb + 10 * nb + 6 * nv + 2 * nv * nv + 2 * nv + 12 * nv + ne * nv + 4 * ne;
  int ints = nc + ne + ws_imax(m.njnt, m.npair) + 4;
  int A = 3 * nb + 4 * nb + ws_imax(ws_imax(6 * m.njnt, 10 * nb), 10 * nc) + (ints + 1) / 2;
  int B = ws_imax(18 * nb + 6 * nv, nv * nv + 3 * ne + ws_quad_doubles(ne, lanes));
  return P + ws_imax(A, B);
}

TK_DEV void ws_init(WS& w, double* base, const DM& m, int ecap, int lanes) {
  int nb = m.nbody, nv = m.nv, ne = ecap, nc = m.maxcontact;
  double* p = base;
  auto take = [&](int n) { double* r = p; p += n; return r; };
  // P
  w.qpos = take(m.nq); w.qvel = take(nv); w.qacc = take(nv); w.warm = take(nv); w.ctrl = take(m.nu);
  w.xipos = take(3 * nb); w.rootcom = take(3 * nb); w.cinert = take(10 * nb); w.cdof = take(6 * nv);
  w.M = take(nv * nv); w.LD = take(nv * nv); w.LDinv = take(nv); w.Hinv = take(nv);
  w.frc_passive = take(nv); w.frc_bias = take(nv); w.frc_act = take(nv); w.frc_smooth = take(nv); w.acc_smooth = take(nv);
  w.frc_con = take(nv); w.Ma = take(nv); w.Mv = take(nv); w.grad = take(nv); w.Mgrad = take(nv); w.search = take(nv);
  w.tmpv = take(nv);
  w.J = take(ne * nv); w.efc_D = take(ne); w.efc_aref = take(ne); w.efc_B = take(ne); w.efc_kterm = take(ne);
  double* u = p;
  // A
  w.xpos = take(3 * nb); w.xquat = take(4 * nb);
  // three tenants of one slot, in program order: joint anchors/axes (kinematics -> com_pos), composite inertias
  // (crb_factor), contacts (collision -> make_constraint)
  double* u2 = p;
  w.xanchor = u2; w.xaxis = u2 + 3 * m.njnt;
  w.crb = u2;
  w.con_dist = u2; w.con_pos = u2 + nc; w.con_frame = u2 + 4 * nc;  // frame: normal + first tangent (6 per contact)
  p = u2 + ws_imax(ws_imax(6 * m.njnt, 10 * nb), 10 * nc);
  // efc_pos / efc_margin only live until the row's impedance is computed: they borrow the slots of efc_aref / efc_kterm
  w.efc_pos = w.efc_aref; w.efc_margin = w.efc_kterm;
  int* ip = reinterpret_cast<int*>(p);
  w.con_pair = ip; ip += nc;
  w.row_info = ip; ip += ne;
  w.cnt = ip;
  // B
  p = u;
  w.cvel = take(6 * nb); w.cacc = take(6 * nb); w.cfrc = take(6 * nb); w.cdof_dot = take(6 * nv);
  p = u;
  w.H = take(nv * nv);
  w.jar = take(ne); w.jv = take(ne); w.efc_force = take(ne); w.quad = take(ws_quad_doubles(ne, lanes));
  w.ncon = 0; w.nefc = 0; w.ecap = ecap; w.work = 0;
#ifdef TOKP_STAGE_CLOCKS
  w.clk = 0; w.clk_on = 0;
#endif
}

}  // namespace tk
