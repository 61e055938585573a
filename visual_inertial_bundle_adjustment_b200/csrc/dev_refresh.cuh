// Per-iteration refresher of the rolling-shutter tables on the device:
// RollingShutterData::compute (lib/motion/preintegration/RollingShutterData.cpp:16-65), reached every
// LM iteration through SingleSessionAdapter::updateRollingShutterData
// (viba/single_session/InitCalibration.cpp:316-325), with forEachIntegratedMeasurement /
// enumIntegrationSteps (PreIntegration.cpp:28-111,308-343), the measurement compensation of
// ImuMeasurementModelParameters::getCompensatedImuMeasurement (ImuMeasurementModelParameters.h:87-101)
// and combine / uncombineLeft / differentiate / integrate of MotionIntegral.cpp:28-41,91-160.
//
// One thread refreshes one rig: ~2 x (half window / IMU period) sequential integration steps over the
// raw IMU-0 samples around the rig's timestamp, with the CURRENT calibration of the rig's window, written
// straight into the table the RS visual kernel reads (layout in dev_factors_ext.cuh).
//
// The file is self-contained and free of device-only constructs so that the test suite can also compile
// it as host code (tests/cpp/rs_refresh_host.cpp) and compare it with the oracle without a GPU; the
// library itself only ever runs it inside rsRefreshKernel.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define VB_RF __host__ __device__ __forceinline__
#else
#define VB_RF inline
#endif

namespace vb {
namespace rf {

constexpr int kMaxSamples = 32;
constexpr int kInterpOff = 1 + kMaxSamples * 11;
constexpr int kStride = kInterpOff + (kMaxSamples - 1) * 9;
constexpr int64_t kMarginNs = 1000;   // PreIntegration.cpp:52

enum { kOk = 0, kNoMargin = 1, kTooManySamples = 2, kNotIncreasing = 3 };

// rotation (unit quaternion x,y,z,w), velocity and position increments over dt: RotVelPos
struct Seg {
  double q[4], dV[3], dP[3], dt;
};

VB_RF void cross(const double* a, const double* b, double* o) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}
VB_RF void quatMul(const double* a, const double* b, double* o) {
  const double x = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
  const double y = a[3] * b[1] + a[1] * b[3] + a[2] * b[0] - a[0] * b[2];
  const double z = a[3] * b[2] + a[2] * b[3] + a[0] * b[1] - a[1] * b[0];
  const double w = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
  o[0] = x; o[1] = y; o[2] = z; o[3] = w;
}
// v' = v + w t + u x t with t = 2 u x v (u = vector part); sgn = -1 rotates by the inverse
VB_RF void quatRot(const double* q, double sgn, const double* v, double* o) {
  const double u[3] = {sgn * q[0], sgn * q[1], sgn * q[2]};
  double t[3], s[3];
  cross(u, v, t);
  t[0] *= 2.0; t[1] *= 2.0; t[2] *= 2.0;
  cross(u, t, s);
  for (int i = 0; i < 3; i++) o[i] = v[i] + q[3] * t[i] + s[i];
}
VB_RF void expQ(const double* w, double* q) {
  const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
  double im, re;
  if (th2 < 1e-20) {
    im = 0.5 - th2 / 48.0 + th2 * th2 / 3840.0;
    re = 1.0 - th2 / 8.0 + th2 * th2 / 384.0;
  } else {
    const double th = sqrt(th2);
    im = sin(0.5 * th) / th;
    re = cos(0.5 * th);
  }
  q[0] = im * w[0]; q[1] = im * w[1]; q[2] = im * w[2]; q[3] = re;
}
VB_RF void logQ(const double* q, double* w) {
  const double n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2];
  double f;
  if (n2 < 1e-20) {
    f = 2.0 / q[3] - (2.0 / 3.0) * n2 / (q[3] * q[3] * q[3]);
  } else {
    const double n = sqrt(n2);
    f = 2.0 * (q[3] < 0 ? atan2(-n, -q[3]) : atan2(n, q[3])) / n;
  }
  w[0] = f * q[0]; w[1] = f * q[1]; w[2] = f * q[2];
}

// MotionIntegral.cpp:117-160 (constant rate and specific force over dt; F6 = 729 as there)
VB_RF void integrate(const double* gyro, const double* accel, double dt, Seg& r) {
  double om[3], up[3], upd[3], a[3], b[3];
  for (int i = 0; i < 3; i++) { om[i] = gyro[i] * dt; up[i] = accel[i] * dt; upd[i] = up[i] * dt; }
  expQ(om, r.q);
  const double th2 = om[0] * om[0] + om[1] * om[1] + om[2] * om[2], th = sqrt(th2), th4 = th2 * th2;
  double c1, c2, c3;
  if (th < 1e-3) {
    c1 = 0.5 - th2 / 24.0 + th4 / 729.0;
    c2 = 1.0 / 6.0 - th2 / 120.0 + th4 / 5040.0;
    c3 = 1.0 / 24.0 - th2 / 729.0 + th4 / 40320.0;
  } else {
    c1 = (1.0 - cos(th)) / th2;
    c2 = (1.0 - sin(th) / th) / th2;
    c3 = (0.5 - c1) / th2;
  }
  cross(om, up, a); cross(om, a, b);
  for (int i = 0; i < 3; i++) r.dV[i] = up[i] + c1 * a[i] + c2 * b[i];
  cross(om, upd, a); cross(om, a, b);
  for (int i = 0; i < 3; i++) r.dP[i] = 0.5 * upd[i] + c2 * a[i] + c3 * b[i];
  r.dt = dt;
}

// c = a then b (MotionIntegral.cpp:28-33); safe for &c == &a
VB_RF void combine(const Seg& a, const Seg& b, Seg& c) {
  double rv[3], rp[3], q[4];
  quatRot(a.q, 1.0, b.dV, rv);
  quatRot(a.q, 1.0, b.dP, rp);
  quatMul(a.q, b.q, q);
  for (int i = 0; i < 3; i++) {
    const double av = a.dV[i];
    c.dP[i] = a.dP[i] + av * b.dt + rp[i];
    c.dV[i] = av + rv[i];
  }
  for (int i = 0; i < 4; i++) c.q[i] = q[i];
  c.dt = a.dt + b.dt;
}

// b with c = combine(a, b) (MotionIntegral.cpp:35-41); safe for &b == &c
VB_RF void uncombineLeft(const Seg& c, const Seg& a, Seg& b) {
  const double dt = c.dt - a.dt;
  double dv[3], dp[3], q[4];
  const double ai[4] = {-a.q[0], -a.q[1], -a.q[2], a.q[3]};
  for (int i = 0; i < 3; i++) {
    dv[i] = c.dV[i] - a.dV[i];
    dp[i] = c.dP[i] - a.dP[i] - a.dV[i] * dt;
  }
  quatMul(ai, c.q, q);
  quatRot(a.q, -1.0, dv, b.dV);
  quatRot(a.q, -1.0, dp, b.dP);
  for (int i = 0; i < 4; i++) b.q[i] = q[i];
  b.dt = dt;
}

// constant rate / specific force / velocity correction reproducing a segment (MotionIntegral.cpp:91-115)
VB_RF void differentiate(const Seg& s, double* interp9) {
  double om[3], ov[3], oov[3];
  logQ(s.q, om);
  const double th2 = om[0] * om[0] + om[1] * om[1] + om[2] * om[2], th = sqrt(th2);
  double q2;
  if (th < 1e-3) {
    q2 = 1.0 / 12.0 - th2 / (4.0 * 180.0) + (th2 * th2) / (16.0 * 1890.0);
  } else {
    const double h = 0.5 * th;
    q2 = (1.0 - h * cos(h) / sin(h)) / th2;
  }
  cross(om, s.dV, ov);
  cross(om, ov, oov);
  double g[3], a[3];
  for (int i = 0; i < 3; i++) {
    g[i] = om[i] / s.dt;
    a[i] = (s.dV[i] - 0.5 * ov[i] + q2 * oov[i]) / s.dt;
  }
  Seg rec;
  integrate(g, a, s.dt, rec);
  for (int i = 0; i < 3; i++) {
    interp9[i] = g[i];
    interp9[3 + i] = a[i];
    interp9[6 + i] = (s.dP[i] - rec.dP[i]) / s.dt;
  }
}

// inverse of a row-major 3x3 by cofactors
VB_RF void inv3(const double* m, double* o) {
  const double c0 = m[4] * m[8] - m[5] * m[7], c1 = m[5] * m[6] - m[3] * m[8], c2 = m[3] * m[7] - m[4] * m[6];
  const double id = 1.0 / (m[0] * c0 + m[1] * c1 + m[2] * c2);
  o[0] = c0 * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
  o[3] = c1 * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
  o[6] = c2 * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}

// measurement model from the 32 doubles of an IMU calibration variable (ImuCalibParam.cpp:214-231:
// gyro scale, accel scale, gyro bias, accel bias, gyro / accel non-orthogonality column-major, accel and
// gyro reference time offsets): compensated = (diag(scale) * nonorth)^-1 * measured - bias
struct ImuModel {
  double gyroInv[9], accelInv[9], gyroBias[3], accelBias[3];
  int64_t dtGyroNs, dtAccelNs;
};
VB_RF void imuModel(const double* c, ImuModel& m) {
  double G[9], A[9];
  for (int r = 0; r < 3; r++)
    for (int k = 0; k < 3; k++) {
      G[r * 3 + k] = c[r] * c[12 + k * 3 + r];
      A[r * 3 + k] = c[3 + r] * c[21 + k * 3 + r];
    }
  inv3(G, m.gyroInv);
  inv3(A, m.accelInv);
  for (int i = 0; i < 3; i++) { m.gyroBias[i] = c[6 + i]; m.accelBias[i] = c[9 + i]; }
  m.dtAccelNs = (int64_t)(c[30] * 1e9);
  m.dtGyroNs = (int64_t)(c[31] * 1e9);
}

// first sample with time > t (PreIntegration.cpp:16-26); n when there is none
VB_RF int64_t firstAfter(const int64_t* timeNs, int64_t n, int64_t t) {
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (t < timeNs[mid]) hi = mid; else lo = mid + 1;
  }
  return lo;
}

// Integrate [startUs, endUs] piece by piece (a piece ends wherever the gyro or the accel sample changes,
// each stream shifted by its own time offset) and append the running integral to the table at every
// gyro boundary, as forEachIntegratedMeasurement + the lambda of RollingShutterData.cpp:25-31 do.
// `acc` is the running integral (identity on entry), `count` the number of table rows so far.
VB_RF int integrateWindow(const int64_t* timeNs, const float* gyro, const float* accel, int64_t n, const ImuModel& m,
                          int64_t startUs, int64_t endUs, Seg& acc, double* rows, int& count) {
  const int64_t t0 = startUs * 1000, t1 = endUs * 1000;
  const int64_t gFirst = firstAfter(timeNs, n, t0 + m.dtGyroNs + kMarginNs), gLast = firstAfter(timeNs, n, t1 + m.dtGyroNs - kMarginNs);
  const int64_t aFirst = firstAfter(timeNs, n, t0 + m.dtAccelNs + kMarginNs), aLast = firstAfter(timeNs, n, t1 + m.dtAccelNs - kMarginNs);
  if (gFirst <= 0 || aFirst <= 0 || gLast >= n || aLast >= n) return kNoMargin;
  auto push = [&](const Seg& s) -> bool {
    if (count >= kMaxSamples) return false;
    double* d = rows + count * 11;
    for (int i = 0; i < 4; i++) d[i] = s.q[i];
    for (int i = 0; i < 3; i++) { d[4 + i] = s.dV[i]; d[7 + i] = s.dP[i]; }
    d[10] = s.dt;
    count++;
    return true;
  };
  int64_t prevStamp = t0;
  int64_t gi = gFirst, ai = aFirst;
  while (gi <= gLast && ai <= aLast) {
    const int64_t gEnd = timeNs[gi] - m.dtGyroNs, aEnd = timeNs[ai] - m.dtAccelNs;
    const int64_t pieceEnd = gEnd < aEnd ? gEnd : aEnd;
    const bool started = gi > gFirst || ai > aFirst;
    const bool newGyro = started && (timeNs[gi - 1] - m.dtGyroNs == prevStamp);
    const bool last = gi >= gLast && ai >= aLast;
    const int64_t stamp = last ? t1 : pieceEnd;
    const double dt = (double)(stamp - prevStamp) * 1e-9;
    prevStamp = stamp;
    if (newGyro || acc.dt == 0.0)
      if (!push(acc)) return kTooManySamples;
    double g[3], a[3];
    for (int r = 0; r < 3; r++) {
      g[r] = m.gyroInv[r * 3] * (double)gyro[gi * 3] + m.gyroInv[r * 3 + 1] * (double)gyro[gi * 3 + 1] +
          m.gyroInv[r * 3 + 2] * (double)gyro[gi * 3 + 2] - m.gyroBias[r];
      a[r] = m.accelInv[r * 3] * (double)accel[ai * 3] + m.accelInv[r * 3 + 1] * (double)accel[ai * 3 + 1] +
          m.accelInv[r * 3 + 2] * (double)accel[ai * 3 + 2] - m.accelBias[r];
    }
    Seg piece;
    integrate(g, a, dt, piece);
    combine(acc, piece, acc);
    gi += (gEnd == pieceEnd);
    ai += (aEnd == pieceEnd);
  }
  return push(acc) ? kOk : kTooManySamples;
}

VB_RF void loadRow(const double* d, Seg& s) {
  for (int i = 0; i < 4; i++) s.q[i] = d[i];
  for (int i = 0; i < 3; i++) { s.dV[i] = d[4 + i]; s.dP[i] = d[7 + i]; }
  s.dt = d[10];
}
VB_RF void storeRow(const Seg& s, double* d) {
  for (int i = 0; i < 4; i++) d[i] = s.q[i];
  for (int i = 0; i < 3; i++) { d[4 + i] = s.dV[i]; d[7 + i] = s.dP[i]; }
  d[10] = s.dt;
}

// RollingShutterData::compute for one rig: table = [n, n x (q, dV, dP, dt), (n-1) x interp]
VB_RF int refreshRig(const int64_t* timeNs, const float* gyro, const float* accel, int64_t n, const double* calib32,
                     int64_t midUs, int64_t halfUs, double* table) {
  ImuModel m;
  imuModel(calib32, m);
  double* rows = table + 1;
  int count = 0;
  Seg acc;
  acc.q[0] = acc.q[1] = acc.q[2] = 0.0; acc.q[3] = 1.0;
  for (int i = 0; i < 3; i++) acc.dV[i] = acc.dP[i] = 0.0;
  acc.dt = 0.0;
  int st = integrateWindow(timeNs, gyro, accel, n, m, midUs - halfUs, midUs, acc, rows, count);
  if (st != kOk) return st;
  // the last row is start -> mid; re-express the earlier ones relative to the midpoint (negative times)
  count--;
  Seg toMid, s;
  loadRow(rows + count * 11, toMid);
  for (int k = 0; k < count; k++) {
    loadRow(rows + k * 11, s);
    uncombineLeft(s, toMid, s);
    storeRow(s, rows + k * 11);
  }
  acc.q[0] = acc.q[1] = acc.q[2] = 0.0; acc.q[3] = 1.0;
  for (int i = 0; i < 3; i++) acc.dV[i] = acc.dP[i] = 0.0;
  acc.dt = 0.0;
  st = integrateWindow(timeNs, gyro, accel, n, m, midUs, midUs + halfUs, acc, rows, count);
  if (st != kOk) return st;
  Seg prev, d;
  loadRow(rows, prev);
  for (int k = 1; k < count; k++) {
    loadRow(rows + k * 11, s);
    if (!(prev.dt < s.dt)) return kNotIncreasing;
    uncombineLeft(s, prev, d);
    differentiate(d, table + kInterpOff + (k - 1) * 9);
    prev = s;
  }
  table[0] = (double)count;
  return kOk;
}

}  // namespace rf
}  // namespace vb
