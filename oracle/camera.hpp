// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// Fisheye624 projection with analytic Jacobians, restating
// deps/projectaria_tools/core/calibration/camera_projections/FisheyeRadTanThinPrism.h:79-239,413-441
// (FisheyeRadTanThinPrism<6, useTangential, useThinPrism, useSingleFocalLength=true>):
// params = [f, cu, cv, k0..k5, p0, p1, s0..s3] (15), and the z < 1e-6 rejection of
// interfaces/ark/camera_model/CameraModelParam.h:35-55.
// Pinned by tests/test_oracle_math.py against the reference's own test recipe
// (CameraProjectionTest.cpp:137-195: analytic vs numeric Jacobians at (1,1,5), tol 1e-4).
#pragma once
#include <cmath>
#include <limits>

namespace orc {

constexpr int kFisheyeNumParams = 15;
constexpr int kStartK = 3, kNumK = 6, kStartP = 9, kStartS = 11;

// returns false when the point is rejected (z < 1e-6).  dPoint: 2x3 row-major, dParam: 2x15
// row-major; either may be null.
inline bool fisheye624Project(const double* p, const double* prm, double* uv, double* dPoint, double* dParam) {
  if (p[2] < 1e-6) return false;
  const double inv_z = 1.0 / p[2];
  const double a = p[0] * inv_z, b = p[1] * inv_z;
  const double a2 = a * a, b2 = b * b;
  const double r_sq = a2 + b2;
  const double r = std::sqrt(r_sq);
  const double th = std::atan(r);
  const double thetaSq = th * th;

  double th_radial = 1.0;
  double theta2is = thetaSq;
  for (int i = 0; i < kNumK; ++i) {
    th_radial += theta2is * prm[kStartK + i];
    theta2is *= thetaSq;
  }
  const double th_divr = (r < std::numeric_limits<double>::epsilon()) ? 1.0 : th / r;
  const double xr = (th_radial * th_divr) * a, yr = (th_radial * th_divr) * b;
  const double rr = xr * xr + yr * yr;

  double ud = xr, vd = yr;
  const double p0 = prm[kStartP], p1 = prm[kStartP + 1];
  {
    const double temp = 2.0 * (xr * p0 + yr * p1);
    ud += temp * xr + rr * p0;
    vd += temp * yr + rr * p1;
  }
  const double rp2 = rr, rp4 = rr * rr;
  ud += prm[kStartS] * rp2 + prm[kStartS + 1] * rp4;
  vd += prm[kStartS + 2] * rp2 + prm[kStartS + 3] * rp4;

  if (dPoint || dParam) {
    // d(uvDistorted)/d(xr,yr)   (FisheyeRadTanThinPrism.h:413-441)
    double d00 = 1.0 + 6.0 * xr * p0 + 2.0 * yr * p1;
    const double offdiag = 2.0 * (xr * p1 + yr * p0);
    double d01 = offdiag, d10 = offdiag;
    double d11 = 1.0 + 6.0 * yr * p1 + 2.0 * xr * p0;
    const double t1 = 2.0 * (prm[kStartS] + 2.0 * prm[kStartS + 1] * rr);
    d00 += xr * t1;
    d01 += yr * t1;
    const double t2 = 2.0 * (prm[kStartS + 2] + 2.0 * prm[kStartS + 3] * rr);
    d10 += xr * t2;
    d11 += yr * t2;

    if (dPoint) {
      double e00, e01, e10, e11; // d(uvDistorted)/d(a,b)
      if (r == 0.0) {
        e00 = 1; e01 = 0; e10 = 0; e11 = 1;
      } else {
        double dthD_dth = 1.0;
        double theta2i = thetaSq;
        for (int i = 0; i < kNumK; ++i) {
          dthD_dth += double(2 * i + 3) * prm[kStartK + i] * theta2i;
          theta2i *= thetaSq;
        }
        const double w1 = dthD_dth / (r_sq + r_sq * r_sq);
        const double w2 = th_radial * th_divr / r_sq;
        const double ab10 = a * b;
        const double m00 = w1 * a2 + w2 * b2;
        const double m01 = (w1 - w2) * ab10;
        const double m11 = w1 * b2 + w2 * a2;
        e00 = d00 * m00 + d01 * m01;
        e01 = d00 * m01 + d01 * m11;
        e10 = d10 * m00 + d11 * m01;
        e11 = d10 * m01 + d11 * m11;
      }
      const double s = prm[0] * inv_z;
      dPoint[0] = s * e00; dPoint[1] = s * e01;
      dPoint[3] = s * e10; dPoint[4] = s * e11;
      dPoint[2] = -(dPoint[0] * a + dPoint[1] * b);
      dPoint[5] = -(dPoint[3] * a + dPoint[4] * b);
    }
    if (dParam) {
      double* r0 = dParam;
      double* r1 = dParam + kFisheyeNumParams;
      for (int i = 0; i < kFisheyeNumParams; i++) r0[i] = r1[i] = 0.0;
      r0[0] = ud; r1[0] = vd;       // d/df
      r0[1] = 1.0; r1[2] = 1.0;     // d/d(cu,cv)
      const double tx = (prm[0] * th_divr) * (d00 * a + d01 * b);
      const double ty = (prm[0] * th_divr) * (d10 * a + d11 * b);
      double theta2i = thetaSq;
      for (int i = 0; i < kNumK; ++i) {
        r0[kStartK + i] = theta2i * tx;
        r1[kStartK + i] = theta2i * ty;
        theta2i *= thetaSq;
      }
      const double f2 = 2.0 * prm[0];
      r0[kStartP] = f2 * xr * xr; r0[kStartP + 1] = f2 * xr * yr;
      r1[kStartP] = f2 * yr * xr; r1[kStartP + 1] = f2 * yr * yr;
      const double t3 = prm[0] * rr;
      r0[kStartP] += t3;
      r1[kStartP + 1] += t3;
      r0[kStartS] = prm[0] * rp2; r0[kStartS + 1] = prm[0] * rp4;
      r1[kStartS + 2] = prm[0] * rp2; r1[kStartS + 3] = prm[0] * rp4;
    }
  }
  uv[0] = prm[0] * ud + prm[1];
  uv[1] = prm[0] * vd + prm[2];
  return true;
}

} // namespace orc
