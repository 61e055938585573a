// ORACLE (test infrastructure, CPU only).  Not part of the product path: only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline/reference legs may use anything in oracle/.
//
// Mini Lie-group / small-matrix library restating the Sophus 1.24.6 conventions the reference
// relies on (Sophus is not vendored in /root/reference: SURVEY.md App. E).  Call sites it serves:
// lib/small_thing/Variable.h:96-127 (SE3 boxPlus/boxMinus), viba/problem/VisualFactor.cpp:48-77,
// viba/problem/InertialFactor.cpp:47-109, viba/problem/RandomWalkFactor.cpp:112-114.
// Parity of the Lie routines themselves is "unpinned" by reference source (dependency absent);
// they are pinned instead by identities in tests/test_oracle_math.py (exp/log round trips,
// numeric Jacobians), as the reference's own tests do for its motion code.
#pragma once
#include <cmath>
#include <cstring>

namespace orc {

// ---- 3-vectors / 3x3 row-major matrices -------------------------------------------------
inline void v3set(double* o, double a, double b, double c) { o[0] = a; o[1] = b; o[2] = c; }
inline double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline void cross3(const double* a, const double* b, double* o) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  o[0] = x; o[1] = y; o[2] = z;
}
// hat(v): o (3x3 row-major)
inline void hat3(const double* v, double* o) {
  o[0] = 0; o[1] = -v[2]; o[2] = v[1];
  o[3] = v[2]; o[4] = 0; o[5] = -v[0];
  o[6] = -v[1]; o[7] = v[0]; o[8] = 0;
}
// C(m x n) = A(m x k) * B(k x n), all row-major
inline void matmul(const double* A, const double* B, double* C, int m, int k, int n) {
  for (int i = 0; i < m; i++)
    for (int j = 0; j < n; j++) {
      double s = 0;
      for (int l = 0; l < k; l++) s += A[i * k + l] * B[l * n + j];
      C[i * n + j] = s;
    }
}
// C(m x n) = A^T(k x m)^T * B(k x n): A is k x m
inline void matmulTN(const double* A, const double* B, double* C, int k, int m, int n) {
  for (int i = 0; i < m; i++)
    for (int j = 0; j < n; j++) {
      double s = 0;
      for (int l = 0; l < k; l++) s += A[l * m + i] * B[l * n + j];
      C[i * n + j] = s;
    }
}
inline void transpose(const double* A, double* B, int m, int n) {
  for (int i = 0; i < m; i++)
    for (int j = 0; j < n; j++) B[j * m + i] = A[i * n + j];
}
inline void mat3vec(const double* M, const double* v, double* o) {
  double x = M[0] * v[0] + M[1] * v[1] + M[2] * v[2];
  double y = M[3] * v[0] + M[4] * v[1] + M[5] * v[2];
  double z = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}
inline void mat3Tvec(const double* M, const double* v, double* o) {
  double x = M[0] * v[0] + M[3] * v[1] + M[6] * v[2];
  double y = M[1] * v[0] + M[4] * v[1] + M[7] * v[2];
  double z = M[2] * v[0] + M[5] * v[1] + M[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}

// ---- SO3 as unit quaternion, memory order (x,y,z,w) like Eigen::Quaterniond::coeffs() -----
inline void quatToMat(const double* q, double* R) {
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w;
  const double txx = tx * x, txy = ty * x, txz = tz * x;
  const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}
// q = a * b
inline void quatMul(const double* a, const double* b, double* o) {
  const double ax = a[0], ay = a[1], az = a[2], aw = a[3];
  const double bx = b[0], by = b[1], bz = b[2], bw = b[3];
  o[0] = aw * bx + ax * bw + ay * bz - az * by;
  o[1] = aw * by + ay * bw + az * bx - ax * bz;
  o[2] = aw * bz + az * bw + ax * by - ay * bx;
  o[3] = aw * bw - ax * bx - ay * by - az * bz;
}
inline void quatConj(const double* q, double* o) { o[0] = -q[0]; o[1] = -q[1]; o[2] = -q[2]; o[3] = q[3]; }
inline void quatNormalize(double* q) {
  double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n;
}
// Sophus SO3::exp (quaternion form, Taylor branch below eps^2 = 1e-20)
inline void so3Exp(const double* w, double* q) {
  const double th2 = dot3(w, w);
  double imag, real;
  if (th2 < 1e-20) {
    const double th4 = th2 * th2;
    imag = 0.5 - th2 / 48.0 + th4 / 3840.0;
    real = 1.0 - th2 / 8.0 + th4 / 384.0;
  } else {
    const double th = std::sqrt(th2), h = 0.5 * th;
    imag = std::sin(h) / th;
    real = std::cos(h);
  }
  q[0] = imag * w[0]; q[1] = imag * w[1]; q[2] = imag * w[2]; q[3] = real;
}
// Sophus SO3::log
inline void so3Log(const double* q, double* w) {
  const double n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2];
  const double qw = q[3];
  double f;
  if (n2 < 1e-20) {
    f = 2.0 / qw - (2.0 / 3.0) * n2 / (qw * qw * qw);
  } else {
    const double n = std::sqrt(n2);
    const double at = (qw < 0) ? std::atan2(-n, -qw) : std::atan2(n, qw);
    f = 2.0 * at / n;
  }
  w[0] = f * q[0]; w[1] = f * q[1]; w[2] = f * q[2];
}
// SO3 left Jacobian  J = I + (1-cos)/th^2 W + (th - sin)/th^3 W^2
inline void so3LeftJacobian(const double* w, double* J) {
  const double th2 = dot3(w, w);
  double W[9], W2[9];
  hat3(w, W);
  matmul(W, W, W2, 3, 3, 3);
  double a, b;
  if (th2 < 1e-20) {
    a = 0.5; b = 1.0 / 6.0;
  } else {
    const double th = std::sqrt(th2);
    a = (1 - std::cos(th)) / th2;
    b = (th - std::sin(th)) / (th2 * th);
  }
  for (int i = 0; i < 9; i++) J[i] = a * W[i] + b * W2[i];
  J[0] += 1; J[4] += 1; J[8] += 1;
}
// SO3 left Jacobian inverse  I - W/2 + (1/th^2 - (1+cos)/(2 th sin)) W^2
inline void so3LeftJacobianInverse(const double* w, double* J) {
  const double th2 = dot3(w, w);
  double W[9], W2[9];
  hat3(w, W);
  matmul(W, W, W2, 3, 3, 3);
  double c;
  if (th2 < 1e-20) {
    c = 1.0 / 12.0;
  } else {
    const double th = std::sqrt(th2);
    c = 1.0 / th2 - (1 + std::cos(th)) / (2 * th * std::sin(th));
  }
  for (int i = 0; i < 9; i++) J[i] = -0.5 * W[i] + c * W2[i];
  J[0] += 1; J[4] += 1; J[8] += 1;
}

// ---- SE3: data = [qx,qy,qz,qw,tx,ty,tz]; tangent = [upsilon(3), omega(3)] ----------------
inline void se3Act(const double* T, const double* p, double* o) {
  double R[9];
  quatToMat(T, R);
  double r[3];
  mat3vec(R, p, r);
  o[0] = r[0] + T[4]; o[1] = r[1] + T[5]; o[2] = r[2] + T[6];
}
inline void se3Mul(const double* A, const double* B, double* O) {
  double q[4], R[9], t[3];
  quatMul(A, B, q);
  quatToMat(A, R);
  mat3vec(R, B + 4, t);
  O[0] = q[0]; O[1] = q[1]; O[2] = q[2]; O[3] = q[3];
  O[4] = t[0] + A[4]; O[5] = t[1] + A[5]; O[6] = t[2] + A[6];
}
inline void se3Inv(const double* T, double* O) {
  double qi[4], R[9], t[3];
  quatConj(T, qi);
  quatToMat(qi, R);
  mat3vec(R, T + 4, t);
  O[0] = qi[0]; O[1] = qi[1]; O[2] = qi[2]; O[3] = qi[3];
  O[4] = -t[0]; O[5] = -t[1]; O[6] = -t[2];
}
inline void se3Exp(const double* d, double* T) {
  const double* u = d;
  const double* w = d + 3;
  so3Exp(w, T);
  double V[9];
  so3LeftJacobian(w, V);
  mat3vec(V, u, T + 4);
}
inline void se3Log(const double* T, double* d) {
  double w[3];
  so3Log(T, w);
  double Vi[9];
  so3LeftJacobianInverse(w, Vi);
  mat3vec(Vi, T + 4, d);
  d[3] = w[0]; d[4] = w[1]; d[5] = w[2];
}
// Adj(T) 6x6 row-major for tangent order [upsilon, omega]: [[R, hat(t) R],[0, R]]
inline void se3Adj(const double* T, double* A) {
  double R[9], H[9], HR[9];
  quatToMat(T, R);
  hat3(T + 4, H);
  matmul(H, R, HR, 3, 3, 3);
  std::memset(A, 0, 36 * sizeof(double));
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      A[i * 6 + j] = R[i * 3 + j];
      A[i * 6 + 3 + j] = HR[i * 3 + j];
      A[(i + 3) * 6 + 3 + j] = R[i * 3 + j];
    }
}
// Barfoot's Q(upsilon, omega) used by the SE3 left Jacobian
inline void se3Q(const double* u, const double* w, double* Q) {
  double U[9], W[9];
  hat3(u, U);
  hat3(w, W);
  const double th2 = dot3(w, w);
  double c1, c2, c3;
  if (th2 < 1e-4) {
    // series (the closed forms cancel catastrophically for small angles)
    const double th4 = th2 * th2;
    c1 = 1.0 / 6.0 - th2 / 120.0 + th4 / 5040.0;
    c2 = 1.0 / 24.0 - th2 / 720.0 + th4 / 40320.0;
    c3 = 1.0 / 120.0 - th2 / 2520.0 + th4 / 120960.0;
  } else {
    const double th = std::sqrt(th2), s = std::sin(th), c = std::cos(th);
    const double th3 = th2 * th, th4 = th2 * th2, th5 = th4 * th;
    c1 = (th - s) / th3;
    c2 = (th2 + 2 * c - 2) / (2 * th4);
    c3 = (2 * th - 3 * s + th * c) / (2 * th5);
  }
  double WU[9], UW[9], WUW[9], WWU[9], UWW[9], WUWW[9], WWUW[9];
  matmul(W, U, WU, 3, 3, 3);
  matmul(U, W, UW, 3, 3, 3);
  matmul(WU, W, WUW, 3, 3, 3);
  matmul(W, WU, WWU, 3, 3, 3);
  matmul(UW, W, UWW, 3, 3, 3);
  matmul(WUW, W, WUWW, 3, 3, 3);
  matmul(W, WUW, WWUW, 3, 3, 3);
  for (int i = 0; i < 9; i++)
    Q[i] = 0.5 * U[i] + c1 * (WU[i] + UW[i] + WUW[i]) + c2 * (WWU[i] + UWW[i] - 3 * WUW[i]) +
        c3 * (WUWW[i] + WWUW[i]);
}
// SE3 left Jacobian inverse 6x6: [[Ji, -Ji Q Ji],[0, Ji]]
inline void se3LeftJacobianInverse(const double* d, double* J) {
  double Ji[9], Q[9], T1[9], T2[9];
  so3LeftJacobianInverse(d + 3, Ji);
  se3Q(d, d + 3, Q);
  matmul(Ji, Q, T1, 3, 3, 3);
  matmul(T1, Ji, T2, 3, 3, 3);
  std::memset(J, 0, 36 * sizeof(double));
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      J[i * 6 + j] = Ji[i * 3 + j];
      J[i * 6 + 3 + j] = -T2[i * 3 + j];
      J[(i + 3) * 6 + 3 + j] = Ji[i * 3 + j];
    }
}

} // namespace orc
