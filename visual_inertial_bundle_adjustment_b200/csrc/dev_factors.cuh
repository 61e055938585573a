// Device residual / Jacobian routines, one per factor type of SURVEY.md App. B.  Each cites the
// reference functor it replaces.  Jacobians are row-major E x d_i; a null J[i] = not requested
// (constant variable or cost-only evaluation; the NullRef convention of Common.h:123-204).
#pragma once
#include "dev_math.cuh"
#include "../../include/viba_b200.h"

namespace vb {

// ---- Fisheye624 (FisheyeRadTanThinPrism<6,true,true,true>) ----------------------------------------
// deps/projectaria_tools/core/calibration/camera_projections/FisheyeRadTanThinPrism.h:79-239,413-441
// + the z < 1e-6 rejection of interfaces/ark/camera_model/CameraModelParam.h:35-55.
// prm = [f, cu, cv, k0..k5, p0, p1, s0..s3].  dP: 2x3, dC: 2x15 (row-major), optional.
template <bool WANT_DP, bool WANT_DC>
VB_DEV bool fisheye624(const double* p, const double* prm, double* uv, double* dP, double* dC) {
  if (p[2] < 1e-6) return false;
  const double iz = 1.0 / p[2];
  const double a = p[0] * iz, b = p[1] * iz;
  const double a2 = a * a, b2 = b * b, r2 = a2 + b2;
  const double r = sqrt(r2);
  const double th = atan(r);
  const double t2 = th * th;
  double rad = 1.0, dRad = 1.0, ti = t2;
#pragma unroll
  for (int i = 0; i < 6; i++) {
    rad = fma(ti, prm[3 + i], rad);
    dRad = fma(double(2 * i + 3) * prm[3 + i], ti, dRad);
    ti *= t2;
  }
  const double thdr = (r < 2.220446049250313e-16) ? 1.0 : th / r;
  const double sc = rad * thdr;
  const double xr = sc * a, yr = sc * b;
  const double rr = xr * xr + yr * yr;
  const double p0 = prm[9], p1 = prm[10];
  const double tmp = 2.0 * (xr * p0 + yr * p1);
  double ud = xr + tmp * xr + rr * p0;
  double vd = yr + tmp * yr + rr * p1;
  const double rr2 = rr * rr;
  ud += prm[11] * rr + prm[12] * rr2;
  vd += prm[13] * rr + prm[14] * rr2;
  const double f = prm[0];
  uv[0] = f * ud + prm[1];
  uv[1] = f * vd + prm[2];
  if (WANT_DP || WANT_DC) {
    double d00 = 1.0 + 6.0 * xr * p0 + 2.0 * yr * p1;
    const double od = 2.0 * (xr * p1 + yr * p0);
    double d01 = od, d10 = od;
    double d11 = 1.0 + 6.0 * yr * p1 + 2.0 * xr * p0;
    const double t1 = 2.0 * (prm[11] + 2.0 * prm[12] * rr);
    d00 += xr * t1; d01 += yr * t1;
    const double t3 = 2.0 * (prm[13] + 2.0 * prm[14] * rr);
    d10 += xr * t3; d11 += yr * t3;
    if (WANT_DP) {
      double e00, e01, e10, e11;
      if (r == 0.0) { e00 = 1; e01 = 0; e10 = 0; e11 = 1; }
      else {
        const double w1 = dRad / (r2 + r2 * r2);
        const double w2 = sc / r2;
        const double ab = a * b;
        const double m00 = w1 * a2 + w2 * b2, m01 = (w1 - w2) * ab, m11 = w1 * b2 + w2 * a2;
        e00 = d00 * m00 + d01 * m01; e01 = d00 * m01 + d01 * m11;
        e10 = d10 * m00 + d11 * m01; e11 = d10 * m01 + d11 * m11;
      }
      const double s = f * iz;
      dP[0] = s * e00; dP[1] = s * e01; dP[3] = s * e10; dP[4] = s * e11;
      dP[2] = -(dP[0] * a + dP[1] * b);
      dP[5] = -(dP[3] * a + dP[4] * b);
    }
    if (WANT_DC) {
#pragma unroll
      for (int i = 0; i < 30; i++) dC[i] = 0.0;
      dC[0] = ud; dC[15] = vd;
      dC[1] = 1.0; dC[15 + 2] = 1.0;
      const double tx = (f * thdr) * (d00 * a + d01 * b);
      const double ty = (f * thdr) * (d10 * a + d11 * b);
      double tj = t2;
#pragma unroll
      for (int i = 0; i < 6; i++) { dC[3 + i] = tj * tx; dC[15 + 3 + i] = tj * ty; tj *= t2; }
      const double f2 = 2.0 * f, fr = f * rr;
      dC[9] = f2 * xr * xr + fr; dC[10] = f2 * xr * yr;
      dC[15 + 9] = f2 * yr * xr; dC[15 + 10] = f2 * yr * yr + fr;
      dC[11] = fr; dC[12] = f * rr2;
      dC[15 + 13] = fr; dC[15 + 14] = f * rr2;
    }
  }
  return true;
}

// ---- VisualFactor::operator() — viba/problem/VisualFactor.cpp:40-82 -----------------------------
// c = [z_u, z_v, sqrtH(2x2)]; Jacobian outputs: JX 2x3, JT 2x6 (rig pose), JE 2x6 (extrinsics),
// JC 2x15 (intrinsics; further columns of a 16/17-dim camera stay zero for this factor).
template <bool WANT_J>
VB_DEV bool visual_factor(const double* X, const double* Tbw, const double* Tcb, const double* cam,
                          const double* c, double* r, double* JX, double* JT, double* JE, double* JC) {
  double Rbw[9], Rcb[9], pr[3], pc[3];
  q2R(Tbw, Rbw);
  q2R(Tcb, Rcb);
  m3v(Rbw, X, pr);
  pr[0] += Tbw[4]; pr[1] += Tbw[5]; pr[2] += Tbw[6];
  m3v(Rcb, pr, pc);
  pc[0] += Tcb[4]; pc[1] += Tcb[5]; pc[2] += Tcb[6];
  double uv[2], dP[6], dC[30];
  if (!fisheye624<WANT_J, WANT_J>(pc, cam, uv, dP, dC)) return false;
  const double e0 = uv[0] - c[0], e1 = uv[1] - c[1];
  r[0] = c[2] * e0 + c[3] * e1;
  r[1] = c[4] * e0 + c[5] * e1;
  if (WANT_J) {
    double W[6]; // sqrtH * dProj/dPointCam
#pragma unroll
    for (int j = 0; j < 3; j++) { W[j] = c[2] * dP[j] + c[3] * dP[3 + j]; W[3 + j] = c[4] * dP[j] + c[5] * dP[3 + j]; }
    double A[6]; // W * Rcb
    mm<2, 3, 3>(W, Rcb, A);
    mm<2, 3, 3>(A, Rbw, JX);
    // [A, A*hat(-pr)] ; row a of A*hat(-p) = p x a  (a x hat(-p))
#pragma unroll
    for (int i = 0; i < 2; i++) {
      const double* a = A + 3 * i;
      JT[i * 6 + 0] = a[0]; JT[i * 6 + 1] = a[1]; JT[i * 6 + 2] = a[2];
      JT[i * 6 + 3] = a[2] * pr[1] - a[1] * pr[2];
      JT[i * 6 + 4] = a[0] * pr[2] - a[2] * pr[0];
      JT[i * 6 + 5] = a[1] * pr[0] - a[0] * pr[1];
      const double* w = W + 3 * i;
      JE[i * 6 + 0] = w[0]; JE[i * 6 + 1] = w[1]; JE[i * 6 + 2] = w[2];
      JE[i * 6 + 3] = w[2] * pc[1] - w[1] * pc[2];
      JE[i * 6 + 4] = w[0] * pc[2] - w[2] * pc[0];
      JE[i * 6 + 5] = w[1] * pc[0] - w[0] * pc[1];
    }
#pragma unroll
    for (int j = 0; j < 15; j++) {
      JC[j] = c[2] * dC[j] + c[3] * dC[15 + j];
      JC[15 + j] = c[4] * dC[j] + c[5] * dC[15 + j];
    }
  }
  return true;
}

// ---- IMU calibration variable (interfaces/ark/imu_model/ImuCalibParam.cpp:55-178,214-244) ---------
// data: gs3 as3 gb3 ab3 gN9(col-major) aN9(col-major) dtA dtG ; flags = ImuCalibrationOptions bits
struct ImuIdx {
  int gb, ab, gs, as, gn, an, rdt, gadt, size;
  VB_DEV explicit ImuIdx(uint32_t f) {
    int r = 0;
    gb = (f & 1) ? r : -1; r += (f & 1) ? 3 : 0;
    ab = (f & 2) ? r : -1; r += (f & 2) ? 3 : 0;
    gs = (f & 4) ? r : -1; r += (f & 4) ? 3 : 0;
    as = (f & 8) ? r : -1; r += (f & 8) ? 3 : 0;
    gn = (f & 16) ? r : -1; r += (f & 16) ? 6 : 0;
    an = (f & 32) ? r : -1; r += (f & 32) ? 3 : 0;
    rdt = (f & 64) ? r : -1; r += (f & 64) ? 1 : 0;
    gadt = (f & 128) ? r : -1; r += (f & 128) ? 1 : 0;
    size = r;
  }
};
// res = c [-] ref (entries of disabled groups untouched)
VB_DEV void imu_boxminus(const double* c, const double* ref, uint32_t flags, double* res) {
  const ImuIdx ix(flags);
  if (ix.gb >= 0) for (int i = 0; i < 3; i++) res[ix.gb + i] = c[6 + i] - ref[6 + i];
  if (ix.ab >= 0) for (int i = 0; i < 3; i++) res[ix.ab + i] = c[9 + i] - ref[9 + i];
  if (ix.gs >= 0) for (int i = 0; i < 3; i++) res[ix.gs + i] = 1.0 / c[i] - 1.0 / ref[i];
  if (ix.as >= 0) for (int i = 0; i < 3; i++) res[ix.as + i] = 1.0 / c[3 + i] - 1.0 / ref[3 + i];
  if (ix.gn >= 0) { // (0,1),(0,2),(1,0),(1,2),(2,0),(2,1) of a column-major 3x3 at offset 12
    const int e[6] = {3, 6, 1, 7, 2, 5};
    for (int i = 0; i < 6; i++) res[ix.gn + i] = c[12 + e[i]] - ref[12 + e[i]];
  }
  if (ix.an >= 0) { // (0,1),(0,2),(1,2) at offset 21
    const int e[3] = {3, 6, 7};
    for (int i = 0; i < 3; i++) res[ix.an + i] = c[21 + e[i]] - ref[21 + e[i]];
  }
  if (ix.rdt >= 0) res[ix.rdt] = c[31] - ref[31];
  if (ix.gadt >= 0) res[ix.gadt] = (c[30] - c[31]) - (ref[30] - ref[31]);
}
VB_DEV void imu_boxplus(double* c, const double* s, uint32_t flags) {
  const ImuIdx ix(flags);
  if (ix.gb >= 0) for (int i = 0; i < 3; i++) c[6 + i] += s[ix.gb + i];
  if (ix.ab >= 0) for (int i = 0; i < 3; i++) c[9 + i] += s[ix.ab + i];
  if (ix.gs >= 0) for (int i = 0; i < 3; i++) { double inv = 1.0 / c[i]; inv += s[ix.gs + i]; c[i] = 1.0 / inv; }
  if (ix.as >= 0) for (int i = 0; i < 3; i++) { double inv = 1.0 / c[3 + i]; inv += s[ix.as + i]; c[3 + i] = 1.0 / inv; }
  if (ix.gn >= 0) {
    double* G = c + 12;
    const double* d = s + ix.gn;
    G[3] += d[0]; G[6] += d[1]; G[1] += d[2]; G[7] += d[3]; G[2] += d[4]; G[5] += d[5];
    G[0] = sqrt(1.0 - (G[3] * G[3] + G[6] * G[6]));
    G[4] = sqrt(1.0 - G[1] * G[1] - G[7] * G[7]);
    G[8] = sqrt(1.0 - (G[2] * G[2] + G[5] * G[5]));
  }
  if (ix.an >= 0) {
    double* A = c + 21;
    const double* d = s + ix.an;
    A[3] += d[0]; A[6] += d[1]; A[7] += d[2];
    A[0] = sqrt(1.0 - (A[3] * A[3] + A[6] * A[6]));
    A[4] = sqrt(1.0 - A[7] * A[7]);
    A[8] = 1.0;
  }
  if (ix.rdt >= 0) { c[31] += s[ix.rdt]; c[30] += s[ix.rdt]; }
  if (ix.gadt >= 0) c[30] += s[ix.gadt];
}
VB_DEV void cam_boxminus(const double* v, const double* base, uint32_t flags, double* d) {
  int n = 15;
  for (int i = 0; i < 15; i++) d[i] = v[i] - base[i];
  if (flags & VB_CAM_EST_READOUT) d[n++] = v[15] - base[15];
  if (flags & VB_CAM_EST_TIMEOFFSET) d[n++] = v[16] - base[16];
}

} // namespace vb
