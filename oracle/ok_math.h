// ORACLE (test infrastructure only).  Scalar-templated restatements of the MuJoCo utility
// routines the pipeline uses (engine_util_blas / engine_util_spatial / engine_util_misc semantics,
// MuJoCo 3.1-series; MuJoCo itself is absent from /root/reference - SURVEY.md F1-F3).
#pragma once
#include <cmath>

#include "ok_count.h"

namespace ok {

constexpr double kMinVal = 1e-15;  // mjMINVAL
constexpr double kPi = 3.14159265358979323846;

template <class R> inline void zero(R* a, int n) { for (int i = 0; i < n; i++) a[i] = R(0.0); }
template <class R> inline void copy(R* a, const R* b, int n) { for (int i = 0; i < n; i++) a[i] = b[i]; }
template <class R> inline R dot3(const R* a, const R* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <class R> inline R dotn(const R* a, const R* b, int n) {
  R s(0.0);
  for (int i = 0; i < n; i++) s += a[i] * b[i];
  return s;
}
template <class R> inline void cross(R* r, const R* a, const R* b) {
  R x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
// mju_normalize3: returns the norm; degenerate vectors become (1,0,0)
template <class R> inline R normalize3(R* a) {
  using std::sqrt;
  R n = sqrt(dot3(a, a));
  if (n < R(kMinVal)) { a[0] = R(1.0); a[1] = R(0.0); a[2] = R(0.0); }
  else { R inv = R(1.0) / n; a[0] *= inv; a[1] *= inv; a[2] *= inv; }
  return n;
}
template <class R> inline void normalize4(R* q) {
  using std::sqrt;
  R n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < R(kMinVal)) { q[0] = R(1.0); q[1] = R(0.0); q[2] = R(0.0); q[3] = R(0.0); }
  else if (fabs(n - R(1.0)) > R(kMinVal)) { R inv = R(1.0) / n; for (int i = 0; i < 4; i++) q[i] *= inv; }
}
template <class R> inline void mul_quat(R* r, const R* a, const R* b) {
  R w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  R x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  R y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  R z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
template <class R> inline void quat2mat(R* m, const R* q) {
  R q00 = q[0] * q[0], q01 = q[0] * q[1], q02 = q[0] * q[2], q03 = q[0] * q[3];
  R q11 = q[1] * q[1], q12 = q[1] * q[2], q13 = q[1] * q[3];
  R q22 = q[2] * q[2], q23 = q[2] * q[3], q33 = q[3] * q[3];
  m[0] = q00 + q11 - q22 - q33; m[4] = q00 - q11 + q22 - q33; m[8] = q00 - q11 - q22 + q33;
  m[1] = R(2.0) * (q12 - q03); m[2] = R(2.0) * (q13 + q02);
  m[3] = R(2.0) * (q12 + q03); m[5] = R(2.0) * (q23 - q01);
  m[6] = R(2.0) * (q13 - q02); m[7] = R(2.0) * (q23 + q01);
}
// mju_rotVecQuat
template <class R> inline void rot_vec_quat(R* r, const R* v, const R* q) {
  if (q[0] == R(1.0) && q[1] == R(0.0) && q[2] == R(0.0) && q[3] == R(0.0)) { r[0] = v[0]; r[1] = v[1]; r[2] = v[2]; return; }
  R m[9];
  quat2mat(m, q);
  R x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  R y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  R z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
template <class R> inline void mul_mat_vec3(R* r, const R* m, const R* v) {
  R x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  R y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  R z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
// mju_axisAngle2Quat
template <class R> inline void axis_angle2quat(R* q, const R* axis, R angle) {
  using std::sin; using std::cos;
  if (angle == R(0.0)) { q[0] = R(1.0); q[1] = R(0.0); q[2] = R(0.0); q[3] = R(0.0); return; }
  R s = sin(angle * R(0.5));
  q[0] = cos(angle * R(0.5));
  q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}
// mju_quatIntegrate: q <- q * exp(vel*scale)
template <class R> inline void quat_integrate(R* q, const R* vel, R scale) {
  R tmp[3] = {vel[0], vel[1], vel[2]};
  R angle = scale * normalize3(tmp);
  R qrot[4], res[4];
  axis_angle2quat(qrot, tmp, angle);
  normalize4(q);
  mul_quat(res, q, qrot);
  for (int i = 0; i < 4; i++) q[i] = res[i];
}
// mju_subQuat: res = 3D velocity taking qa to qb  (qb * conj(qa) expressed in qa's frame)
template <class R> inline void sub_quat(R* res, const R* qa, const R* qb) {
  using std::atan2;
  R qneg[4] = {qa[0], -qa[1], -qa[2], -qa[3]}, qdif[4];
  mul_quat(qdif, qneg, qb);
  R axis[3] = {qdif[1], qdif[2], qdif[3]};
  R sin_a_2 = normalize3(axis);
  R speed = R(2.0) * atan2(sin_a_2, qdif[0]);
  if (speed > R(kPi)) speed -= R(2.0 * kPi);
  res[0] = axis[0] * speed; res[1] = axis[1] * speed; res[2] = axis[2] * speed;
}

// ---- spatial algebra: vectors are [rot(3); lin(3)], 10-number inertias ------------------------
// mju_mulInertVec
template <class R> inline void mul_inert_vec(R* res, const R* i, const R* v) {
  res[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
  res[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
  res[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
  res[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
  res[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
  res[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
}
// mju_inertCom: inertia (principal `inert`, orientation `mat`) about a point offset by -dif from the com
template <class R> inline void inert_com(R* res, const R* inert, const R* mat, const R* dif, R mass) {
  R tmp[9];
  // tmp = diag(inert) * mat'
  tmp[0] = mat[0] * inert[0]; tmp[3] = mat[1] * inert[1]; tmp[6] = mat[2] * inert[2];
  tmp[1] = mat[3] * inert[0]; tmp[4] = mat[4] * inert[1]; tmp[7] = mat[5] * inert[2];
  tmp[2] = mat[6] * inert[0]; tmp[5] = mat[7] * inert[1]; tmp[8] = mat[8] * inert[2];
  // res_rot = mat * diag(inert) * mat'
  res[0] = mat[0] * tmp[0] + mat[1] * tmp[3] + mat[2] * tmp[6];
  res[1] = mat[3] * tmp[1] + mat[4] * tmp[4] + mat[5] * tmp[7];
  res[2] = mat[6] * tmp[2] + mat[7] * tmp[5] + mat[8] * tmp[8];
  res[3] = mat[0] * tmp[1] + mat[1] * tmp[4] + mat[2] * tmp[7];
  res[4] = mat[0] * tmp[2] + mat[1] * tmp[5] + mat[2] * tmp[8];
  res[5] = mat[3] * tmp[2] + mat[4] * tmp[5] + mat[5] * tmp[8];
  // res_rot -= mass * dif_cross * dif_cross
  res[0] += mass * (dif[1] * dif[1] + dif[2] * dif[2]);
  res[1] += mass * (dif[0] * dif[0] + dif[2] * dif[2]);
  res[2] += mass * (dif[0] * dif[0] + dif[1] * dif[1]);
  res[3] -= mass * dif[0] * dif[1];
  res[4] -= mass * dif[0] * dif[2];
  res[5] -= mass * dif[1] * dif[2];
  // res_tran = mass * dif ; res_mass = mass
  res[6] = mass * dif[0]; res[7] = mass * dif[1]; res[8] = mass * dif[2];
  res[9] = mass;
}
// mju_dofCom: 6D motion axis of a rotation about `axis` through a point at -offset from the com
template <class R> inline void dof_com(R* res, const R* axis, const R* offset) {
  res[0] = axis[0]; res[1] = axis[1]; res[2] = axis[2];
  if (offset) cross(res + 3, axis, offset);
  else { res[3] = R(0.0); res[4] = R(0.0); res[5] = R(0.0); }
}
// mju_crossMotion
template <class R> inline void cross_motion(R* res, const R* vel, const R* v) {
  R a[3], b[3], c[3];
  cross(a, vel, v);
  cross(b, vel, v + 3);
  cross(c, vel + 3, v);
  res[0] = a[0]; res[1] = a[1]; res[2] = a[2];
  res[3] = b[0] + c[0]; res[4] = b[1] + c[1]; res[5] = b[2] + c[2];
}
// mju_crossForce
template <class R> inline void cross_force(R* res, const R* vel, const R* f) {
  R a[3], b[3], c[3];
  cross(a, vel, f);
  cross(b, vel + 3, f + 3);
  cross(c, vel, f + 3);
  res[0] = a[0] + b[0]; res[1] = a[1] + b[1]; res[2] = a[2] + b[2];
  res[3] = c[0]; res[4] = c[1]; res[5] = c[2];
}

// mju_makeFrame: complete frame[0:3] (normal) to a right-handed orthonormal frame
template <class R> inline void make_frame(R* frame) {
  using std::sqrt;
  normalize3(frame);
  R yn = sqrt(dot3(frame + 3, frame + 3));
  if (yn < R(0.5)) {
    frame[3] = R(0.0); frame[4] = R(0.0); frame[5] = R(0.0);
    if (frame[1] < R(0.5) && frame[1] > R(-0.5)) frame[4] = R(1.0);
    else frame[5] = R(1.0);
  }
  R tmp = dot3(frame, frame + 3);
  frame[3] -= tmp * frame[0]; frame[4] -= tmp * frame[1]; frame[5] -= tmp * frame[2];
  normalize3(frame + 3);
  cross(frame + 6, frame, frame + 3);
}

}  // namespace ok
