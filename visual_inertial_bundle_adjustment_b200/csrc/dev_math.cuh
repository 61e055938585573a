// Device-side small-matrix / Lie-group routines for the factor kernels (sm_100a, FP64).
// Conventions follow what the reference gets from Sophus 1.24.6 / Eigen (not vendored, SURVEY.md
// App. E): quaternion memory order (x,y,z,w), SE3 data = [q, t], SE3 tangent = [upsilon, omega],
// left-multiplicative retraction exp(d)*T (lib/small_thing/Variable.h:96-127).
// Everything is register-resident: fixed-size arrays with compile-time loop bounds.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace vb {

#define VB_DEV __device__ __forceinline__

VB_DEV double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
VB_DEV void cross3(const double* a, const double* b, double* o) {
  const double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  o[0] = x; o[1] = y; o[2] = z;
}
VB_DEV void hat3(const double* v, double* o) {
  o[0] = 0; o[1] = -v[2]; o[2] = v[1];
  o[3] = v[2]; o[4] = 0; o[5] = -v[0];
  o[6] = -v[1]; o[7] = v[0]; o[8] = 0;
}
template <int M, int K, int N>
VB_DEV void mm(const double* A, const double* B, double* C) { // C = A(MxK) B(KxN), row-major
#pragma unroll
  for (int i = 0; i < M; i++)
#pragma unroll
    for (int j = 0; j < N; j++) {
      double s = 0;
#pragma unroll
      for (int l = 0; l < K; l++) s = fma(A[i * K + l], B[l * N + j], s);
      C[i * N + j] = s;
    }
}
VB_DEV void m3v(const double* M, const double* v, double* o) {
  const double x = M[0] * v[0] + M[1] * v[1] + M[2] * v[2];
  const double y = M[3] * v[0] + M[4] * v[1] + M[5] * v[2];
  const double z = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}
VB_DEV void m3tv(const double* M, const double* v, double* o) {
  const double x = M[0] * v[0] + M[3] * v[1] + M[6] * v[2];
  const double y = M[1] * v[0] + M[4] * v[1] + M[7] * v[2];
  const double z = M[2] * v[0] + M[5] * v[1] + M[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}

// ---- unit quaternions -----------------------------------------------------------------------
VB_DEV void q2R(const double* q, double* R) {
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w;
  const double txx = tx * x, txy = ty * x, txz = tz * x;
  const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}
VB_DEV void qmul(const double* a, const double* b, double* o) {
  const double ax = a[0], ay = a[1], az = a[2], aw = a[3];
  const double bx = b[0], by = b[1], bz = b[2], bw = b[3];
  o[0] = aw * bx + ax * bw + ay * bz - az * by;
  o[1] = aw * by + ay * bw + az * bx - ax * bz;
  o[2] = aw * bz + az * bw + ax * by - ay * bx;
  o[3] = aw * bw - ax * bx - ay * by - az * bz;
}
VB_DEV void qconj(const double* q, double* o) { o[0] = -q[0]; o[1] = -q[1]; o[2] = -q[2]; o[3] = q[3]; }

VB_DEV void so3_exp(const double* w, double* q) {
  const double th2 = dot3(w, w);
  double im, re;
  if (th2 < 1e-20) {
    const double th4 = th2 * th2;
    im = 0.5 - th2 / 48.0 + th4 / 3840.0;
    re = 1.0 - th2 / 8.0 + th4 / 384.0;
  } else {
    const double th = sqrt(th2);
    double s, c;
    sincos(0.5 * th, &s, &c);
    im = s / th; re = c;
  }
  q[0] = im * w[0]; q[1] = im * w[1]; q[2] = im * w[2]; q[3] = re;
}
VB_DEV void so3_log(const double* q, double* w) {
  const double n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2];
  const double qw = q[3];
  double f;
  if (n2 < 1e-20) {
    f = 2.0 / qw - (2.0 / 3.0) * n2 / (qw * qw * qw);
  } else {
    const double n = sqrt(n2);
    const double at = (qw < 0) ? atan2(-n, -qw) : atan2(n, qw);
    f = 2.0 * at / n;
  }
  w[0] = f * q[0]; w[1] = f * q[1]; w[2] = f * q[2];
}
// J = I + a W + b W^2 with W = hat(w); W^2 = w w^T - |w|^2 I
VB_DEV void so3_axpby(const double* w, double a, double b, double* J) {
  const double th2 = dot3(w, w);
  J[0] = 1 + b * (w[0] * w[0] - th2); J[1] = -a * w[2] + b * w[0] * w[1]; J[2] = a * w[1] + b * w[0] * w[2];
  J[3] = a * w[2] + b * w[1] * w[0]; J[4] = 1 + b * (w[1] * w[1] - th2); J[5] = -a * w[0] + b * w[1] * w[2];
  J[6] = -a * w[1] + b * w[2] * w[0]; J[7] = a * w[0] + b * w[2] * w[1]; J[8] = 1 + b * (w[2] * w[2] - th2);
}
VB_DEV void so3_left_jac(const double* w, double* J) {
  const double th2 = dot3(w, w);
  double a, b;
  if (th2 < 1e-20) { a = 0.5; b = 1.0 / 6.0; }
  else {
    const double th = sqrt(th2);
    double s, c;
    sincos(th, &s, &c);
    a = (1 - c) / th2; b = (th - s) / (th2 * th);
  }
  so3_axpby(w, a, b, J);
}
VB_DEV void so3_left_jac_inv(const double* w, double* J) {
  const double th2 = dot3(w, w);
  double c2;
  if (th2 < 1e-20) c2 = 1.0 / 12.0;
  else {
    const double th = sqrt(th2);
    double s, c;
    sincos(th, &s, &c);
    c2 = 1.0 / th2 - (1 + c) / (2 * th * s);
  }
  so3_axpby(w, -0.5, c2, J);
}

// ---- SE3 -------------------------------------------------------------------------------------
VB_DEV void se3_act(const double* T, const double* p, double* o) {
  double R[9], r[3];
  q2R(T, R);
  m3v(R, p, r);
  o[0] = r[0] + T[4]; o[1] = r[1] + T[5]; o[2] = r[2] + T[6];
}
VB_DEV void se3_mul(const double* A, const double* B, double* O) {
  double q[4], R[9], t[3];
  qmul(A, B, q);
  q2R(A, R);
  m3v(R, B + 4, t);
  O[0] = q[0]; O[1] = q[1]; O[2] = q[2]; O[3] = q[3];
  O[4] = t[0] + A[4]; O[5] = t[1] + A[5]; O[6] = t[2] + A[6];
}
VB_DEV void se3_inv(const double* T, double* O) {
  double qi[4], R[9], t[3];
  qconj(T, qi);
  q2R(qi, R);
  m3v(R, T + 4, t);
  O[0] = qi[0]; O[1] = qi[1]; O[2] = qi[2]; O[3] = qi[3];
  O[4] = -t[0]; O[5] = -t[1]; O[6] = -t[2];
}
VB_DEV void se3_exp(const double* d, double* T) {
  so3_exp(d + 3, T);
  double V[9];
  so3_left_jac(d + 3, V);
  m3v(V, d, T + 4);
}
VB_DEV void se3_log(const double* T, double* d) {
  double w[3], Vi[9];
  so3_log(T, w);
  so3_left_jac_inv(w, Vi);
  m3v(Vi, T + 4, d);
  d[3] = w[0]; d[4] = w[1]; d[5] = w[2];
}
// Adj(T) = [[R, hat(t) R],[0, R]]  (6x6 row-major)
VB_DEV void se3_adj(const double* T, double* A) {
  double R[9], H[9], HR[9];
  q2R(T, R);
  hat3(T + 4, H);
  mm<3, 3, 3>(H, R, HR);
#pragma unroll
  for (int i = 0; i < 36; i++) A[i] = 0;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      A[i * 6 + j] = R[i * 3 + j];
      A[i * 6 + 3 + j] = HR[i * 3 + j];
      A[(i + 3) * 6 + 3 + j] = R[i * 3 + j];
    }
}
// SE3 left Jacobian inverse [[Ji, -Ji Q Ji],[0, Ji]] with Barfoot's Q(upsilon, omega)
VB_DEV void se3_left_jac_inv(const double* d, double* J) {
  const double* u = d;
  const double* w = d + 3;
  double U[9], W[9];
  hat3(u, U);
  hat3(w, W);
  const double th2 = dot3(w, w);
  double c1, c2, c3;
  if (th2 < 1e-4) {
    const double th4 = th2 * th2;
    c1 = 1.0 / 6.0 - th2 / 120.0 + th4 / 5040.0;
    c2 = 1.0 / 24.0 - th2 / 720.0 + th4 / 40320.0;
    c3 = 1.0 / 120.0 - th2 / 2520.0 + th4 / 120960.0;
  } else {
    const double th = sqrt(th2);
    double s, c;
    sincos(th, &s, &c);
    const double th3 = th2 * th, th4 = th2 * th2, th5 = th4 * th;
    c1 = (th - s) / th3;
    c2 = (th2 + 2 * c - 2) / (2 * th4);
    c3 = (2 * th - 3 * s + th * c) / (2 * th5);
  }
  double WU[9], UW[9], WUW[9], WWU[9], UWW[9], WUWW[9], WWUW[9], Q[9];
  mm<3, 3, 3>(W, U, WU);
  mm<3, 3, 3>(U, W, UW);
  mm<3, 3, 3>(WU, W, WUW);
  mm<3, 3, 3>(W, WU, WWU);
  mm<3, 3, 3>(UW, W, UWW);
  mm<3, 3, 3>(WUW, W, WUWW);
  mm<3, 3, 3>(W, WUW, WWUW);
#pragma unroll
  for (int i = 0; i < 9; i++)
    Q[i] = 0.5 * U[i] + c1 * (WU[i] + UW[i] + WUW[i]) + c2 * (WWU[i] + UWW[i] - 3 * WUW[i]) + c3 * (WUWW[i] + WWUW[i]);
  double Ji[9], T1[9], T2[9];
  so3_left_jac_inv(w, Ji);
  mm<3, 3, 3>(Ji, Q, T1);
  mm<3, 3, 3>(T1, Ji, T2);
#pragma unroll
  for (int i = 0; i < 36; i++) J[i] = 0;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      J[i * 6 + j] = Ji[i * 3 + j];
      J[i * 6 + 3 + j] = -T2[i * 3 + j];
      J[(i + 3) * 6 + 3 + j] = Ji[i * 3 + j];
    }
}

// S2::ortho (lib/small_thing/Variable.h:168-180): 2x3 row-major tangent basis
VB_DEV void s2_ortho(const double* v, double* res) {
  double t1[3] = {0, 0, 0};
  const double xa = fabs(v[0]), ya = fabs(v[1]), za = fabs(v[2]);
  const int coord = xa < fmin(ya, za) ? 0 : (ya < za ? 1 : 2);
  t1[coord] = 1.0;
  const double vsq = dot3(v, v), vn = sqrt(vsq);
  const double c = dot3(t1, v) / vsq;
  double r0[3] = {t1[0] - c * v[0], t1[1] - c * v[1], t1[2] - c * v[2]};
  const double n0 = sqrt(dot3(r0, r0));
#pragma unroll
  for (int i = 0; i < 3; i++) r0[i] /= n0;
  double r1[3];
  cross3(r0, v, r1);
#pragma unroll
  for (int i = 0; i < 3; i++) { res[i] = r0[i]; res[3 + i] = r1[i] / vn; }
}

// HuberLossWithCutoff::jet2 (lib/small_thing/SoftLoss.h:131-163); a<=0 or inf => identity
VB_DEV void loss_jet2(double a, double k, double s, double& val, double& der) {
  if (!(a > 0) || isinf(a)) { val = s; der = 1.0; return; }
  const double b = a * a, k2 = k * k;
  if (s > k2) { val = 2.0 * a * k - b; der = 0.0; }
  else if (s > b) { const double r = sqrt(s); val = 2.0 * a * r - b; der = a / r; }
  else { val = s; der = 1.0; }
}

// D(8x8) += A(8x4, row) * B(4x8, col).  Lane l holds a = A[l/4][l%4], b = B[l%4][l/4],
// c = C[l/4][2*(l%4) + {0,1}].
VB_DEV void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// ---- asynchronous global -> shared staging (LDGSTS): every copy is issued without waiting for the
// previous one, so a tile costs about one memory round trip instead of one per element -------------
VB_DEV void cp_async8(double* smemDst, const double* gsrc, bool valid) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smemDst);
  const int sz = valid ? 8 : 0; // src-size 0 => zero fill, nothing is read
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(s), "l"(gsrc), "r"(sz) : "memory");
}
// 16-byte variant (L2 only): srcBytes in {0, 8, 16}, the rest of the 16 bytes is zero filled
VB_DEV void cp_async16(double* smemDst, const double* gsrc, int srcBytes) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smemDst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gsrc), "r"(srcBytes) : "memory");
}
VB_DEV void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
// dst[rowsPad x colsPad] (leading dimension ld) <- src[rows x cols] (row stride srcStride), zero padded.
// Warps take rows, lanes take columns.  Call cp_async_wait_all() + __syncthreads() before use.
VB_DEV void loadTileAsync(double* dst, int ld, int rowsPad, int colsPad, const double* src, int64_t srcStride, int rows, int cols) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  // 16-byte copies when every row of source and destination is 16-byte aligned (half the LDGSTS count)
  if ((((uintptr_t)src | (uintptr_t)dst) & 15) == 0 && ((srcStride | ld | colsPad) & 1) == 0) {
    for (int r = warp; r < rowsPad; r += nw) {
      const bool rv = r < rows;
      const double* srow = src + (rv ? (int64_t)r * srcStride : 0);
      double* drow = dst + r * ld;
      for (int c = 2 * lane; c < colsPad; c += 64) {
        const int nv = rv ? min(2, max(0, cols - c)) : 0;
        cp_async16(drow + c, nv ? srow + c : src, 8 * nv);
      }
    }
    return;
  }
  for (int r = warp; r < rowsPad; r += nw) {
    const bool rv = r < rows;
    const double* srow = src + (rv ? (int64_t)r * srcStride : 0);
    double* drow = dst + r * ld;
    for (int c = lane; c < colsPad; c += 32) {
      const bool v = rv && c < cols;
      cp_async8(drow + c, v ? srow + c : src, v);
    }
  }
}
// global[rows x cols] (row stride dstStride) <- smem tile
VB_DEV void storeTile(double* dst, int64_t dstStride, const double* srcSm, int ld, int rows, int cols) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int r = warp; r < rows; r += nw)
    for (int c = lane; c < cols; c += 32) dst[(int64_t)r * dstStride + c] = srcSm[r * ld + c];
}

// fp64 reduction without return value (REDG.E.ADD.F64)
VB_DEV void red_add(double* addr, double v) { atomicAdd(addr, v); }

} // namespace vb
