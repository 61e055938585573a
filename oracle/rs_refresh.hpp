// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// Per-iteration refresher of the rolling-shutter tables, restated from
//   RollingShutterData::compute                  lib/motion/preintegration/RollingShutterData.cpp:16-65
//   forEachIntegratedMeasurement                 lib/motion/preintegration/PreIntegration.cpp:308-343
//   enumIntegrationSteps / measIndex_GT          PreIntegration.cpp:16-111
//   getCompensatedImuMeasurement                 lib/motion/imu_types/ImuMeasurementModelParameters.h:87-101
//   uncombineLeft / differentiate                lib/motion/preintegration/MotionIntegral.cpp:35-41,91-115
// as called every LM iteration by SingleSessionAdapter::updateRollingShutterData
// (viba/single_session/InitCalibration.cpp:316-325).  The measurement model is read from the 32 doubles
// of the IMU calibration variable (imu.hpp layout).  Parity of this file with the reference is pinned only
// by properties (tests/test_rs_refresh.py: agreement with the closed-form motion of the synthetic
// trajectory, combine/uncombine round trips); the reference's own tests for this code
// (lib/motion/preintegration/tests) need Eigen/Sophus and cannot be built here.
#pragma once
#include <algorithm>
#include <cstdint>
#include <functional>
#include <stdexcept>
#include <vector>
#include "imu.hpp"
#include "motion.hpp"

namespace orc {

struct ImuSamples {           // std::vector<ImuMeasurement>: one time stamp, gyro and accel as float
  std::vector<int64_t> timeNs;
  std::vector<float> gyro, accel;   // 3 per sample
};

struct MeasModel {            // ImuMeasurementModelParameters, the part compensation needs
  double gyroScaleInv[9], accelScaleInv[9], gyroBias[3], accelBias[3];
  double dtRefGyroSec, dtRefAccelSec;
};

inline void inverse3(const double* M, double* O) {
  const double a = M[0], b = M[1], c = M[2], d = M[3], e = M[4], f = M[5], g = M[6], h = M[7], i = M[8];
  const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
  O[0] = (e * i - f * h) / det; O[1] = (c * h - b * i) / det; O[2] = (b * f - c * e) / det;
  O[3] = (f * g - d * i) / det; O[4] = (a * i - c * g) / det; O[5] = (c * d - a * f) / det;
  O[6] = (d * h - e * g) / det; O[7] = (b * g - a * h) / det; O[8] = (a * e - b * d) / det;
}

inline MeasModel measModelFromCalib(const double* c) {
  MeasModel m;
  double G[9], A[9];
  for (int r = 0; r < 3; r++)
    for (int k = 0; k < 3; k++) {   // scale matrix = diag(scaleVec) * nonorth (ImuMeasurementModelParameters.h:121-131)
      G[r * 3 + k] = c[C_GS + r] * c[C_GN + cm(r, k)];
      A[r * 3 + k] = c[C_AS + r] * c[C_AN + cm(r, k)];
    }
  inverse3(G, m.gyroScaleInv);
  inverse3(A, m.accelScaleInv);
  for (int i = 0; i < 3; i++) { m.gyroBias[i] = c[C_GB + i]; m.accelBias[i] = c[C_AB + i]; }
  m.dtRefGyroSec = c[C_DTG];
  m.dtRefAccelSec = c[C_DTA];
  return m;
}

// MotionIntegral.cpp:35-41
inline Rvp uncombineLeft(const Rvp& c, const Rvp& a) {
  Rvp b;
  double ainv[4], t[3], u[3];
  quatConj(a.q, ainv);
  quatMul(ainv, c.q, b.q);
  for (int i = 0; i < 3; i++) t[i] = c.dV[i] - a.dV[i];
  quatRot(ainv, t, b.dV);
  b.dt = c.dt - a.dt;
  for (int i = 0; i < 3; i++) u[i] = c.dP[i] - a.dP[i] - a.dV[i] * b.dt;
  quatRot(ainv, u, b.dP);
  return b;
}

// MotionIntegral.cpp:91-115 -> interp = gyro(3), accel(3), deltaVel(3)
inline void differentiate(const Rvp& rvp, double* interp) {
  double omega[3];
  so3Log(rvp.q, omega);
  const double th2 = dot3(omega, omega), th = std::sqrt(th2);
  const double q1 = -0.5;
  double q2;
  if (th < 1e-3) {
    q2 = 1.0 / 12.0 - th2 / (4.0 * 180.0) + (th2 * th2) / (16.0 * 1890.0);
  } else {
    const double h = th * 0.5;
    q2 = (1.0 - h * std::cos(h) / std::sin(h)) / th2;
  }
  double omegaVel[3], oov[3], gyro[3], accel[3];
  cross3(omega, rvp.dV, omegaVel);
  cross3(omega, omegaVel, oov);
  for (int i = 0; i < 3; i++) {
    gyro[i] = omega[i] / rvp.dt;
    accel[i] = (rvp.dV[i] + q1 * omegaVel[i] + q2 * oov[i]) / rvp.dt;
  }
  const Rvp recon = integrate(gyro, accel, rvp.dt);
  for (int i = 0; i < 3; i++) {
    interp[i] = gyro[i];
    interp[3 + i] = accel[i];
    interp[6 + i] = (rvp.dP[i] - recon.dP[i]) / rvp.dt;
  }
}

// PreIntegration.cpp:16-26
inline int64_t measIndexGT(const ImuSamples& meas, int64_t timeNs) {
  const auto it = std::upper_bound(meas.timeNs.begin(), meas.timeNs.end(), timeNs);   // first sample strictly later
  if (it == meas.timeNs.end()) throw std::runtime_error("measIndex_GT: unexpected, it == meas.end()");
  return (int64_t)(it - meas.timeNs.begin());
}

// PreIntegration.cpp:28-111: fn(gyro sample, accel sample, dtSec, newAccelMeas, newGyroMeas)
inline void enumIntegrationSteps(const ImuSamples& meas, const MeasModel& model, int64_t timeStartUs, int64_t timeEndUs,
                                 const std::function<void(const float*, const float*, double, bool, bool)>& fn) {
  const int64_t dtGyroNs = (int64_t)(model.dtRefGyroSec * 1e9), dtAccelNs = (int64_t)(model.dtRefAccelSec * 1e9);
  const int64_t refStartNs = timeStartUs * 1000, refEndNs = timeEndUs * 1000;
  const int64_t kMarginNs = 1000;
  const int64_t gyroStart = measIndexGT(meas, refStartNs + dtGyroNs + kMarginNs);
  const int64_t gyroEnd = measIndexGT(meas, refEndNs + dtGyroNs - kMarginNs);
  if (gyroStart <= 0) throw std::runtime_error("enumIntegrationSteps: gyro index, not enough margin at beginning of interval");
  const int64_t accelStart = measIndexGT(meas, refStartNs + dtAccelNs + kMarginNs);
  const int64_t accelEnd = measIndexGT(meas, refEndNs + dtAccelNs - kMarginNs);
  if (accelStart <= 0) throw std::runtime_error("enumIntegrationSteps: accel index, not enough margin at beginning of interval");
  int64_t prevStamp = refStartNs;
  for (int64_t gi = gyroStart, ai = accelStart; gi <= gyroEnd && ai <= accelEnd;) {
    const int64_t gNow = gi, aNow = ai;
    const int64_t adjGyro = meas.timeNs[gNow] - dtGyroNs, adjAccel = meas.timeNs[aNow] - dtAccelNs;
    const int64_t endMeas = std::min(adjGyro, adjAccel);
    const bool moved = gNow > gyroStart || aNow > accelStart;
    const bool toNewAccel = moved && (meas.timeNs[aNow - 1] - dtAccelNs == prevStamp);
    const bool toNewGyro = moved && (meas.timeNs[gNow - 1] - dtGyroNs == prevStamp);
    const int64_t endStamp = (gNow >= gyroEnd && aNow >= accelEnd) ? refEndNs : endMeas;
    const double dtSec = (endStamp - prevStamp) * 1e-9;
    prevStamp = endStamp;
    gi += (adjGyro == endMeas);
    ai += (adjAccel == endMeas);
    fn(&meas.gyro[3 * gNow], &meas.accel[3 * aNow], dtSec, toNewAccel, toNewGyro);
  }
}

// PreIntegration.cpp:308-343
inline void forEachIntegratedMeasurement(const ImuSamples& meas, const MeasModel& model, int64_t timeStartUs, int64_t timeEndUs,
                                         const std::function<void(const Rvp&, bool, bool)>& fn) {
  Rvp prev{{0, 0, 0, 1}, {0, 0, 0}, {0, 0, 0}, 0.0};
  enumIntegrationSteps(meas, model, timeStartUs, timeEndUs, [&](const float* g, const float* a, double dtSec, bool newAccel, bool newGyro) {
    const bool atStart = prev.dt == 0.0;
    fn(prev, newAccel || atStart, newGyro || atStart);
    double gd[3] = {g[0], g[1], g[2]}, ad[3] = {a[0], a[1], a[2]}, gc[3], ac[3];
    mat3vec(model.gyroScaleInv, gd, gc);
    mat3vec(model.accelScaleInv, ad, ac);
    for (int i = 0; i < 3; i++) { gc[i] -= model.gyroBias[i]; ac[i] -= model.accelBias[i]; }
    prev = combine(prev, integrate(gc, ac, dtSec));
  });
  fn(prev, true, true);
}

// PreIntegration.cpp:277-306: the preintegrated rotation / velocity / position over [start, end]
inline Rvp integrateMeasurements(const ImuSamples& meas, const MeasModel& model, int64_t timeStartUs, int64_t timeEndUs) {
  Rvp total{{0, 0, 0, 1}, {0, 0, 0}, {0, 0, 0}, 0.0};
  bool first = true;
  enumIntegrationSteps(meas, model, timeStartUs, timeEndUs, [&](const float* g, const float* a, double dtSec, bool, bool) {
    double gd[3] = {g[0], g[1], g[2]}, ad[3] = {a[0], a[1], a[2]}, gc[3], ac[3];
    mat3vec(model.gyroScaleInv, gd, gc);
    mat3vec(model.accelScaleInv, ad, ac);
    for (int i = 0; i < 3; i++) { gc[i] -= model.gyroBias[i]; ac[i] -= model.accelBias[i]; }
    const Rvp piece = integrate(gc, ac, dtSec);
    total = first ? piece : combine(total, piece);
    first = false;
  });
  return total;
}

// RollingShutterData::compute -> the per-rig table of motion.hpp
inline void rsCompute(const ImuSamples& meas, const double* calib32, int64_t midpointUs, int64_t halfLengthUs, double* table) {
  const MeasModel model = measModelFromCalib(calib32);
  std::vector<Rvp> sampled;
  auto keepGyroBoundaries = [&](const Rvp& rvp, bool, bool atGyroBoundary) {
    if (atGyroBoundary) sampled.push_back(rvp);
  };
  forEachIntegratedMeasurement(meas, model, midpointUs - halfLengthUs, midpointUs, keepGyroBoundaries);
  const Rvp startToMid = sampled.back();
  sampled.pop_back();
  for (Rvp& s : sampled) s = uncombineLeft(s, startToMid);   // relative to the midpoint, negative times
  forEachIntegratedMeasurement(meas, model, midpointUs, midpointUs + halfLengthUs, keepGyroBoundaries);
  if ((int)sampled.size() > kRsMaxSamples) throw std::runtime_error("RollingShutterData::compute: more than 32 samples in a window");
  table[0] = (double)sampled.size();
  for (size_t i = 0; i < sampled.size(); i++) {
    double* d = table + 1 + i * 11;
    for (int k = 0; k < 4; k++) d[k] = sampled[i].q[k];
    for (int k = 0; k < 3; k++) { d[4 + k] = sampled[i].dV[k]; d[7 + k] = sampled[i].dP[k]; }
    d[10] = sampled[i].dt;
  }
  for (size_t i = 1; i < sampled.size(); i++) {
    if (sampled[i - 1].dt >= sampled[i].dt) throw std::runtime_error("Wut?");
    differentiate(uncombineLeft(sampled[i], sampled[i - 1]), table + kRsInterpOff + (i - 1) * 9);
  }
}

}  // namespace orc
