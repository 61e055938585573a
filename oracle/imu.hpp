// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// IMU calibration variable: data layout of VarSpec<ImuCalibParam>::getData
// (interfaces/ark/imu_model/ImuCalibParam.cpp:214-231): gyroScale3, accelScale3, gyroBias3,
// accelBias3, gyroNonorth9 (column-major), accelNonorth9 (column-major), dtRefAccel, dtRefGyro.
// Error-state layout: lib/motion/imu_types/ImuCalibrationJacobianIndices.h:37-96.
// boxPlus / boxMinus: interfaces/ark/imu_model/ImuCalibParam.cpp:55-116,119-178.
#pragma once
#include <cmath>
#include <cstdint>

namespace orc {

// var_flags bits for VB_VAR_IMUCALIB = ImuCalibrationOptions
enum : uint32_t {
  IMU_GYRO_BIAS = 1, IMU_ACCEL_BIAS = 2, IMU_GYRO_SCALE = 4, IMU_ACCEL_SCALE = 8,
  IMU_GYRO_NONORTH = 16, IMU_ACCEL_NONORTH = 32, IMU_REF_DT = 64, IMU_GYRO_ACCEL_DT = 128
};

struct ImuIdx {
  int gyroBias = -1, accelBias = -1, gyroScale = -1, accelScale = -1, gyroNonorth = -1,
      accelNonorth = -1, refDt = -1, gyroAccelDt = -1, size = 0;
  explicit ImuIdx(uint32_t f) {
    int r = 0;
    if (f & IMU_GYRO_BIAS) { gyroBias = r; r += 3; }
    if (f & IMU_ACCEL_BIAS) { accelBias = r; r += 3; }
    if (f & IMU_GYRO_SCALE) { gyroScale = r; r += 3; }
    if (f & IMU_ACCEL_SCALE) { accelScale = r; r += 3; }
    if (f & IMU_GYRO_NONORTH) { gyroNonorth = r; r += 6; }
    if (f & IMU_ACCEL_NONORTH) { accelNonorth = r; r += 3; }
    if (f & IMU_REF_DT) { refDt = r; r += 1; }
    if (f & IMU_GYRO_ACCEL_DT) { gyroAccelDt = r; r += 1; }
    size = r;
  }
};

// data offsets
constexpr int C_GS = 0, C_AS = 3, C_GB = 6, C_AB = 9, C_GN = 12, C_AN = 21, C_DTA = 30, C_DTG = 31;
// column-major 3x3 element (r,c)
inline int cm(int r, int c) { return c * 3 + r; }

inline void imuCalibBoxPlus(double* c, const double* s, uint32_t flags) {
  const ImuIdx ix(flags);
  if (ix.gyroBias >= 0) for (int i = 0; i < 3; i++) c[C_GB + i] += s[ix.gyroBias + i];
  if (ix.accelBias >= 0) for (int i = 0; i < 3; i++) c[C_AB + i] += s[ix.accelBias + i];
  if (ix.gyroScale >= 0)
    for (int i = 0; i < 3; i++) { double inv = 1.0 / c[C_GS + i]; inv += s[ix.gyroScale + i]; c[C_GS + i] = 1.0 / inv; }
  if (ix.accelScale >= 0)
    for (int i = 0; i < 3; i++) { double inv = 1.0 / c[C_AS + i]; inv += s[ix.accelScale + i]; c[C_AS + i] = 1.0 / inv; }
  if (ix.gyroNonorth >= 0) {
    double* G = c + C_GN;
    const double* d = s + ix.gyroNonorth;
    G[cm(0, 1)] += d[0]; G[cm(0, 2)] += d[1];
    G[cm(1, 0)] += d[2]; G[cm(1, 2)] += d[3];
    G[cm(2, 0)] += d[4]; G[cm(2, 1)] += d[5];
    G[cm(0, 0)] = std::sqrt(1.0 - (G[cm(0, 1)] * G[cm(0, 1)] + G[cm(0, 2)] * G[cm(0, 2)]));
    G[cm(1, 1)] = std::sqrt(1.0 - G[cm(1, 0)] * G[cm(1, 0)] - G[cm(1, 2)] * G[cm(1, 2)]);
    G[cm(2, 2)] = std::sqrt(1.0 - (G[cm(2, 0)] * G[cm(2, 0)] + G[cm(2, 1)] * G[cm(2, 1)]));
  }
  if (ix.accelNonorth >= 0) {
    double* A = c + C_AN;
    const double* d = s + ix.accelNonorth;
    A[cm(0, 1)] += d[0]; A[cm(0, 2)] += d[1];
    A[cm(1, 2)] += d[2];
    A[cm(0, 0)] = std::sqrt(1.0 - (A[cm(0, 1)] * A[cm(0, 1)] + A[cm(0, 2)] * A[cm(0, 2)]));
    A[cm(1, 1)] = std::sqrt(1.0 - A[cm(1, 2)] * A[cm(1, 2)]);
    A[cm(2, 2)] = 1.0;
  }
  if (ix.refDt >= 0) { c[C_DTG] += s[ix.refDt]; c[C_DTA] += s[ix.refDt]; }
  if (ix.gyroAccelDt >= 0) c[C_DTA] += s[ix.gyroAccelDt];
}

// res = c [-] ref ; entries of res not covered by enabled options are left untouched
inline void imuCalibBoxMinus(const double* c, const double* ref, uint32_t flags, double* res) {
  const ImuIdx ix(flags);
  if (ix.gyroBias >= 0) for (int i = 0; i < 3; i++) res[ix.gyroBias + i] = c[C_GB + i] - ref[C_GB + i];
  if (ix.accelBias >= 0) for (int i = 0; i < 3; i++) res[ix.accelBias + i] = c[C_AB + i] - ref[C_AB + i];
  if (ix.gyroScale >= 0) for (int i = 0; i < 3; i++) res[ix.gyroScale + i] = 1.0 / c[C_GS + i] - 1.0 / ref[C_GS + i];
  if (ix.accelScale >= 0) for (int i = 0; i < 3; i++) res[ix.accelScale + i] = 1.0 / c[C_AS + i] - 1.0 / ref[C_AS + i];
  if (ix.gyroNonorth >= 0) {
    const double *G = c + C_GN, *H = ref + C_GN;
    double* d = res + ix.gyroNonorth;
    d[0] = G[cm(0, 1)] - H[cm(0, 1)]; d[1] = G[cm(0, 2)] - H[cm(0, 2)];
    d[2] = G[cm(1, 0)] - H[cm(1, 0)]; d[3] = G[cm(1, 2)] - H[cm(1, 2)];
    d[4] = G[cm(2, 0)] - H[cm(2, 0)]; d[5] = G[cm(2, 1)] - H[cm(2, 1)];
  }
  if (ix.accelNonorth >= 0) {
    const double *A = c + C_AN, *B = ref + C_AN;
    double* d = res + ix.accelNonorth;
    d[0] = A[cm(0, 1)] - B[cm(0, 1)]; d[1] = A[cm(0, 2)] - B[cm(0, 2)];
    d[2] = A[cm(1, 2)] - B[cm(1, 2)];
  }
  if (ix.refDt >= 0) res[ix.refDt] = c[C_DTG] - ref[C_DTG];
  if (ix.gyroAccelDt >= 0) res[ix.gyroAccelDt] = (c[C_DTA] - c[C_DTG]) - (ref[C_DTA] - ref[C_DTG]);
}

} // namespace orc
