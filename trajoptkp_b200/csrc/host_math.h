// Small host-side vector / quaternion helpers for the model compiler.
#pragma once
#include <cmath>

namespace tokp {
namespace hm {

inline void cross(double r[3], const double a[3], const double b[3]) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
inline double dot3(const double a[3], const double b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline double norm3(const double a[3]) { return std::sqrt(dot3(a, a)); }
inline double normalize3(double a[3]) {
  double n = norm3(a);
  if (n < 1e-15) { a[0] = 1; a[1] = 0; a[2] = 0; return n; }
  a[0] /= n; a[1] /= n; a[2] /= n;
  return n;
}
inline void quat_mul(double r[4], const double a[4], const double b[4]) {
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
inline void quat_normalize(double q[4]) {
  double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < 1e-15) { q[0] = 1; q[1] = q[2] = q[3] = 0; return; }
  for (int i = 0; i < 4; i++) q[i] /= n;
}
// row-major 3x3 rotation from unit quaternion
inline void quat2mat(double m[9], const double q[4]) {
  double q00 = q[0] * q[0], q11 = q[1] * q[1], q22 = q[2] * q[2], q33 = q[3] * q[3];
  double q01 = q[0] * q[1], q02 = q[0] * q[2], q03 = q[0] * q[3];
  double q12 = q[1] * q[2], q13 = q[1] * q[3], q23 = q[2] * q[3];
  m[0] = q00 + q11 - q22 - q33; m[4] = q00 - q11 + q22 - q33; m[8] = q00 - q11 - q22 + q33;
  m[1] = 2 * (q12 - q03); m[2] = 2 * (q13 + q02);
  m[3] = 2 * (q12 + q03); m[5] = 2 * (q23 - q01);
  m[6] = 2 * (q13 - q02); m[7] = 2 * (q23 + q01);
}
inline void rot_vec_quat(double r[3], const double v[3], const double q[4]) {
  double m[9];
  quat2mat(m, q);
  double x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  double y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  double z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
inline void axis_angle2quat(double q[4], const double axis[3], double angle) {
  if (angle == 0) { q[0] = 1; q[1] = q[2] = q[3] = 0; return; }
  double s = std::sin(angle * 0.5);
  q[0] = std::cos(angle * 0.5);
  q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}
// minimal rotation taking +z to `vec` (need not be unit)
inline void z2quat(double q[4], const double vec_in[3]) {
  double vec[3] = {vec_in[0], vec_in[1], vec_in[2]};
  double z[3] = {0, 0, 1};
  q[0] = 1; q[1] = q[2] = q[3] = 0;
  if (normalize3(vec) < 1e-15) return;
  double axis[3];
  cross(axis, z, vec);
  double s = norm3(axis);
  if (s < 1e-10) {
    // parallel or anti-parallel
    if (vec[2] < 0) { q[0] = 0; q[1] = 1; q[2] = 0; q[3] = 0; }
    return;
  }
  axis[0] /= s; axis[1] /= s; axis[2] /= s;
  double ang = std::atan2(s, vec[2]);
  axis_angle2quat(q, axis, ang);
}
// quaternion from row-major rotation matrix
inline void mat2quat(double q[4], const double m[9]) {
  double tr = m[0] + m[4] + m[8];
  if (tr > 0) {
    double s = std::sqrt(tr + 1.0) * 2;
    q[0] = 0.25 * s; q[1] = (m[7] - m[5]) / s; q[2] = (m[2] - m[6]) / s; q[3] = (m[3] - m[1]) / s;
  } else if (m[0] > m[4] && m[0] > m[8]) {
    double s = std::sqrt(1.0 + m[0] - m[4] - m[8]) * 2;
    q[0] = (m[7] - m[5]) / s; q[1] = 0.25 * s; q[2] = (m[1] + m[3]) / s; q[3] = (m[2] + m[6]) / s;
  } else if (m[4] > m[8]) {
    double s = std::sqrt(1.0 + m[4] - m[0] - m[8]) * 2;
    q[0] = (m[2] - m[6]) / s; q[1] = (m[1] + m[3]) / s; q[2] = 0.25 * s; q[3] = (m[5] + m[7]) / s;
  } else {
    double s = std::sqrt(1.0 + m[8] - m[0] - m[4]) * 2;
    q[0] = (m[3] - m[1]) / s; q[1] = (m[2] + m[6]) / s; q[2] = (m[5] + m[7]) / s; q[3] = 0.25 * s;
  }
  quat_normalize(q);
}

// Symmetric 3x3 eigen-decomposition by cyclic Jacobi.  A is row-major symmetric.
// Outputs eigenvalues (descending) and a proper rotation whose COLUMNS are the eigenvectors.
inline void eig3(const double A[9], double eval[3], double evec[9]) {
  double a[3][3] = {{A[0], A[1], A[2]}, {A[3], A[4], A[5]}, {A[6], A[7], A[8]}};
  double v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  for (int sweep = 0; sweep < 64; sweep++) {
    double off = std::fabs(a[0][1]) + std::fabs(a[0][2]) + std::fabs(a[1][2]);
    double diag = std::fabs(a[0][0]) + std::fabs(a[1][1]) + std::fabs(a[2][2]);
    if (off <= 1e-18 * diag || off == 0) break;
    for (int p = 0; p < 2; p++)
      for (int q = p + 1; q < 3; q++) {
        if (a[p][q] == 0) continue;
        double theta = (a[q][q] - a[p][p]) / (2 * a[p][q]);
        double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1));
        double c = 1 / std::sqrt(t * t + 1), s = t * c;
        for (int k = 0; k < 3; k++) {
          double akp = a[k][p], akq = a[k][q];
          a[k][p] = c * akp - s * akq;
          a[k][q] = s * akp + c * akq;
        }
        for (int k = 0; k < 3; k++) {
          double apk = a[p][k], aqk = a[q][k];
          a[p][k] = c * apk - s * aqk;
          a[q][k] = s * apk + c * aqk;
        }
        for (int k = 0; k < 3; k++) {
          double vkp = v[k][p], vkq = v[k][q];
          v[k][p] = c * vkp - s * vkq;
          v[k][q] = s * vkp + c * vkq;
        }
      }
  }
  int idx[3] = {0, 1, 2};
  double e[3] = {a[0][0], a[1][1], a[2][2]};
  for (int i = 0; i < 2; i++)
    for (int j = i + 1; j < 3; j++)
      if (e[idx[j]] > e[idx[i]]) { int t = idx[i]; idx[i] = idx[j]; idx[j] = t; }
  for (int c = 0; c < 3; c++) {
    eval[c] = e[idx[c]];
    for (int r = 0; r < 3; r++) evec[3 * r + c] = v[r][idx[c]];
  }
  // make it a proper rotation (det = +1)
  double c0[3] = {evec[0], evec[3], evec[6]}, c1[3] = {evec[1], evec[4], evec[7]}, c2[3];
  cross(c2, c0, c1);
  evec[2] = c2[0]; evec[5] = c2[1]; evec[8] = c2[2];
}

}  // namespace hm
}  // namespace tokp
