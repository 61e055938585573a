// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// Per-iteration refresher of the preintegration records (second half), restated from
//   computePreIntegration                        lib/motion/preintegration/PreIntegration.cpp:136-275
//   dRvp_dLeftCompensatedMeas / dStartTime / dEndTime                PreIntegration.cpp:113-134
//   getCompensatedImuMeasurementAndJac           lib/motion/preintegration/CompensateJac.cpp:146-249
//   integrate(gyro, accel, dt, &J) / combineJacs lib/motion/preintegration/MotionIntegral.cpp:52-75,162-226
// as driven every LM iteration under --recompute-preint by
//   SingleSessionAdapter::regenerateAllPreintegrationsFromImuMeasurements  viba/single_session/InertialFactors.cpp:19-70
// through the preStepCallback of interfaces/ark/main_AriaKit_ViBa.cpp:95-101.
// The record written is the factor store's aux layout (kPreintStride = 250 doubles):
//   R(4, xyzw) dV(3) dP(3) dt(1) J(9 x 23 row-major, the first E columns used) evalPoint(32)
// plus the 9x9 covariance (its inverse is the factor's information matrix, fixed at factor creation:
// viba/problem/InertialFactor.cpp:313) and omegaAtEnd (captured by the omega priors at creation).
// Pinned with the reference's own test recipes restated in tests/test_preint.py:
// lib/motion/preintegration/tests/TestPreIntegration.cpp:104-203, TestCompensateJac.cpp:91-163,
// TestMotionIntegral.cpp:54-161 (central differences with the reference's steps and tolerances).
#pragma once
#include <cstring>
#include "rs_refresh.hpp"

namespace orc {

constexpr int kPreintRec = 250;
constexpr int kPreintJ = 11, kPreintEval = 218;

// C[m x n] = A[m x k] B[k x n], row-major with explicit leading dimensions
inline void gemmLd(const double* A, int lda, const double* B, int ldb, double* C, int ldc, int m, int k, int n) {
  for (int i = 0; i < m; i++)
    for (int j = 0; j < n; j++) {
      double s = 0;
      for (int t = 0; t < k; t++) s += A[i * lda + t] * B[t * ldb + j];
      C[i * ldc + j] = s;
    }
}

// MotionIntegral.cpp:162-226: rvp = integrate(gyro, accel, dt), J = d rvp / d (gyro, accel)  (9 x 6, row-major)
inline Rvp integrateJac(const double* gyro, const double* accel, double dt, double* J) {
  double om[3], up[3];
  for (int i = 0; i < 3; i++) { om[i] = gyro[i] * dt; up[i] = accel[i] * dt; }
  Rvp r;
  so3Exp(om, r.q);
  const double th2 = dot3(om, om), th = std::sqrt(th2), th4 = th2 * th2;
  double c1, c2, c3, d1, d2, d3;
  if (th < 1e-3) {
    c1 = (1.0 / 2.0) - (th2 / 24.0) + (th4 / 729.0); // F6 = 729 (sic), MotionIntegral.cpp:82
    c2 = (1.0 / 6.0) - (th2 / 120.0) + (th4 / 5040.0);
    c3 = (1.0 / 24.0) - (th2 / 729.0) + (th4 / 40320.0);
    d1 = -(2.0 / 24.0) + th2 * (4.0 / 729.0) + th4 * (6.0 / 40320.0);
    d2 = -(2.0 / 120.0) + th2 * (4.0 / 5040.0) + th4 * (6.0 / 362880.0);
    d3 = -(2.0 / 729.0) + th2 * (4.0 / 40320.0) + th4 * (6.0 / 3628800.0);
  } else {
    const double s = std::sin(th) / th, m = (1.0 - std::cos(th)) / th2;
    c1 = m; c2 = (1.0 - s) / th2; c3 = (0.5 - m) / th2;
    d1 = (s - 2.0 * m) / th2;
    d2 = (m - 3.0 * c2) / th2;
    d3 = (-1.0 - s + 4.0 * m) / th4;
  }
  double W[9], W2[9], U2V[9], U2P[9];
  hat3(om, W);
  matmul(W, W, W2, 3, 3, 3);
  for (int i = 0; i < 9; i++) { U2V[i] = c1 * W[i] + c2 * W2[i]; U2P[i] = c2 * W[i] + c3 * W2[i]; }
  U2V[0] += 1; U2V[4] += 1; U2V[8] += 1;
  U2P[0] += 0.5; U2P[4] += 0.5; U2P[8] += 0.5;
  mat3vec(U2V, up, r.dV);
  const double upd[3] = {up[0] * dt, up[1] * dt, up[2] * dt};
  mat3vec(U2P, upd, r.dP);
  r.dt = dt;
  // Jacobian
  for (int i = 0; i < 54; i++) J[i] = 0.0;
  double nup[3] = {-up[0], -up[1], -up[2]}, DwXu[9], oxu[3], noxu[3], H2[9], WD[9], DwXwXu[9];
  hat3(nup, DwXu);
  for (int i = 0; i < 9; i++) DwXu[i] *= dt;                     // hat(-upsilon) dt
  cross3(om, up, oxu);
  for (int i = 0; i < 3; i++) noxu[i] = -oxu[i];
  hat3(noxu, H2);
  matmul(W, DwXu, WD, 3, 3, 3);
  for (int i = 0; i < 9; i++) DwXwXu[i] = H2[i] * dt + WD[i];    // hat(-(omega x upsilon)) dt + Omega DwXu
  double M1[9], M2[9], VD1[3], PD1[3];
  for (int i = 0; i < 9; i++) { M1[i] = d1 * W[i] + d2 * W2[i]; M2[i] = d2 * W[i] + d3 * W2[i]; }
  mat3vec(M1, up, VD1);
  mat3vec(M2, upd, PD1);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      J[i * 6 + j] = U2V[i * 3 + j] * dt;                                                         // block (0,0): as in the reference
      J[(3 + i) * 6 + j] = VD1[i] * om[j] * dt + c1 * DwXu[i * 3 + j] + c2 * DwXwXu[i * 3 + j];   // JV + JV2
      J[(6 + i) * 6 + j] = PD1[i] * om[j] * dt + (c2 * DwXu[i * 3 + j] + c3 * DwXwXu[i * 3 + j]) * dt; // JP + JP2
      J[(3 + i) * 6 + 3 + j] = U2V[i * 3 + j] * dt;
      J[(6 + i) * 6 + 3 + j] = U2P[i * 3 + j] * dt * dt;
    }
  return r;
}

// MotionIntegral.cpp:52-75: c = combine(a, b) and cJac from aJac, bJac (9 x n each, row-major)
inline Rvp combineJacs(const Rvp& a, const Rvp& b, const double* aJ, const double* bJ, double* cJ, int n) {
  double aR[9], aRbV[3], aRbP[3];
  quatToMat(a.q, aR);
  mat3vec(aR, b.dV, aRbV);
  mat3vec(aR, b.dP, aRbP);
  Rvp c;
  quatMul(a.q, b.q, c.q);
  for (int i = 0; i < 3; i++) { c.dV[i] = a.dV[i] + aRbV[i]; c.dP[i] = a.dP[i] + a.dV[i] * b.dt + aRbP[i]; }
  c.dt = a.dt + b.dt;
  double nV[3] = {-aRbV[0], -aRbV[1], -aRbV[2]}, nP[3] = {-aRbP[0], -aRbP[1], -aRbP[2]}, HV[9], HP[9];
  hat3(nV, HV);
  hat3(nP, HP);
  for (int col = 0; col < n; col++) {
    double at[3], av[3], ap[3], bt[3], bv[3], bp[3];
    for (int i = 0; i < 3; i++) {
      at[i] = aJ[i * n + col]; av[i] = aJ[(3 + i) * n + col]; ap[i] = aJ[(6 + i) * n + col];
      bt[i] = bJ[i * n + col]; bv[i] = bJ[(3 + i) * n + col]; bp[i] = bJ[(6 + i) * n + col];
    }
    double Rt[3], Rv[3], Rp[3], hv[3], hp[3];
    mat3vec(aR, bt, Rt); mat3vec(aR, bv, Rv); mat3vec(aR, bp, Rp);
    mat3vec(HV, at, hv); mat3vec(HP, at, hp);
    for (int i = 0; i < 3; i++) {
      cJ[i * n + col] = at[i] + Rt[i];
      cJ[(3 + i) * n + col] = av[i] + hv[i] + Rv[i];
      cJ[(6 + i) * n + col] = ap[i] + av[i] * b.dt + hp[i] + Rp[i];
    }
  }
  return c;
}

struct CalibMats {   // what getCompensatedImuMeasurementAndJac derives from the parameters (CompensateJac.cpp:160-166)
  double gyroNonorthInv[9], accelNonorthInv[9], gyroScaleMat[9], accelScaleMat[9];
  double gN[9], aN[9], gyroBias[3], accelBias[3];   // nonorth matrices row-major
};
inline CalibMats calibMats(const double* c) {
  CalibMats m;
  for (int r = 0; r < 3; r++)
    for (int k = 0; k < 3; k++) { m.gN[r * 3 + k] = c[C_GN + cm(r, k)]; m.aN[r * 3 + k] = c[C_AN + cm(r, k)]; }
  inverse3(m.gN, m.gyroNonorthInv);
  inverse3(m.aN, m.accelNonorthInv);
  for (int r = 0; r < 3; r++)
    for (int k = 0; k < 3; k++) {   // nonorthInv * diag(1 / scale)
      m.gyroScaleMat[r * 3 + k] = m.gyroNonorthInv[r * 3 + k] / c[C_GS + k];
      m.accelScaleMat[r * 3 + k] = m.accelNonorthInv[r * 3 + k] / c[C_AS + k];
    }
  for (int i = 0; i < 3; i++) { m.gyroBias[i] = c[C_GB + i]; m.accelBias[i] = c[C_AB + i]; }
  return m;
}

// CompensateJac.cpp:146-249.  rawGyro / rawAccel = averageSignal of the step.  calibJac: 6 x E, measJac: 6 x 6; both are
// written into one row-major 6 x (6 + E) matrix cmJ = [measJac | calibJac] (the reference's compMeasJac).
inline void compensateAndJac(const CalibMats& m, const ImuIdx& ix, const double* rawGyro, const double* rawAccel, double* gyro, double* accel, double* cmJ) {
  const int E = ix.size, n = 6 + E;
  for (int i = 0; i < 6 * n; i++) cmJ[i] = 0.0;
  double* cal = cmJ + 6;   // calibJac(r, c) = cal[r * n + c]
  double scaledGyro[3], scaledAccel[3];
  mat3vec(m.gyroScaleMat, rawGyro, scaledGyro);
  mat3vec(m.accelScaleMat, rawAccel, scaledAccel);
  if (ix.gyroScale >= 0)
    for (int r = 0; r < 3; r++) for (int k = 0; k < 3; k++) cal[r * n + ix.gyroScale + k] = m.gyroNonorthInv[r * 3 + k] * rawGyro[k];
  if (ix.gyroNonorth >= 0) {
    static const int kRows[6] = {0, 0, 1, 1, 2, 2}, kCols[6] = {1, 2, 0, 2, 0, 1};
    for (int i = 0; i < 6; i++) {
      const int r = kRows[i], c = kCols[i];
      const double dNrr = -m.gN[r * 3 + c] / m.gN[r * 3 + r];
      const double f = scaledGyro[r] * dNrr + scaledGyro[c];
      for (int k = 0; k < 3; k++) cal[k * n + ix.gyroNonorth + i] = -m.gyroNonorthInv[k * 3 + r] * f;
    }
  }
  for (int i = 0; i < 3; i++) gyro[i] = scaledGyro[i] - m.gyroBias[i];
  for (int r = 0; r < 3; r++) for (int k = 0; k < 3; k++) cmJ[r * n + k] = m.gyroScaleMat[r * 3 + k];
  if (ix.gyroBias >= 0) for (int i = 0; i < 3; i++) cal[i * n + ix.gyroBias + i] = -1.0;
  if (ix.accelScale >= 0)
    for (int r = 0; r < 3; r++) for (int k = 0; k < 3; k++) cal[(3 + r) * n + ix.accelScale + k] = m.accelNonorthInv[r * 3 + k] * rawAccel[k];
  for (int r = 0; r < 3; r++) for (int k = 0; k < 3; k++) cmJ[(3 + r) * n + 3 + k] = m.accelScaleMat[r * 3 + k];
  if (ix.accelNonorth >= 0) {
    static const int kRows[3] = {0, 0, 1}, kCols[3] = {1, 2, 2};
    for (int i = 0; i < 3; i++) {
      const int r = kRows[i], c = kCols[i];
      const double dNrr = -m.aN[r * 3 + c] / m.aN[r * 3 + r];
      const double f = scaledAccel[r] * dNrr + scaledAccel[c];   // scaledAccelAdd_dNonOrth2 is zero in the reference
      for (int k = 0; k < 3; k++) cal[(3 + k) * n + ix.accelNonorth + i] = -m.accelNonorthInv[k * 3 + r] * f;
    }
  }
  for (int i = 0; i < 3; i++) accel[i] = scaledAccel[i] - m.accelBias[i];
  if (ix.accelBias >= 0) for (int i = 0; i < 3; i++) cal[(3 + i) * n + ix.accelBias + i] = -1.0;
}

// PreIntegration.cpp:113-120
inline void dRvpDLeft(const Rvp& rvp, const double* g, const double* a, double* out9) {
  double nV[3] = {-rvp.dV[0], -rvp.dV[1], -rvp.dV[2]}, nP[3] = {-rvp.dP[0], -rvp.dP[1], -rvp.dP[2]}, H[9], t[3];
  for (int i = 0; i < 3; i++) out9[i] = g[i];
  hat3(nV, H); mat3vec(H, g, t);
  for (int i = 0; i < 3; i++) out9[3 + i] = t[i] + a[i];
  hat3(nP, H); mat3vec(H, g, t);
  for (int i = 0; i < 3; i++) out9[6 + i] = a[i] * rvp.dt + t[i];
}

struct PreintResult {
  double rec[kPreintRec];   // the factor store's aux record
  double cov[81];           // rvpCov, row-major
  double omegaAtEnd[3];
  int steps;
};

// PreIntegration.cpp:136-275.  gyroVar / accelVar = ImuNoiseModelParameters::{gyroSampleVarianceRad2Sec2, accelSampleVarianceM2Sec4}.
// Q7 (reference quirk, kept): the gyro-accel time-offset column is written when `jacInd.gyroAccelTimeOffsetIdx()` is
// NON-ZERO (an int used as a condition, :197) — so with the option disabled (index -1) column 15 - 1 = 14, the raw
// accel-z column of the step Jacobian, is overwritten, and with the option as the only one (index 0) it is skipped.
inline PreintResult preintCompute(const ImuSamples& meas, const double* calib32, uint32_t flags, const double* gyroVar, const double* accelVar,
                                  int64_t timeStartUs, int64_t timeEndUs) {
  const ImuIdx ix(flags);
  const int E = ix.size, n = 15 + E, nc = 6 + E;
  const MeasModel model = measModelFromCalib(calib32);   // time offsets for enumIntegrationSteps
  const CalibMats cmats = calibMats(calib32);
  std::vector<double> rvpJac(9 * n, 0.0), rvp1Jac(9 * n, 0.0), rvp2Jac(9 * n, 0.0), cmJ(6 * nc);
  double cov[81] = {0}, fromA[27] = {0}, fromG[27] = {0};
  Rvp total{{0, 0, 0, 1}, {0, 0, 0}, {0, 0, 0}, 0.0};
  bool have = false;
  double startG[3] = {0, 0, 0}, startA[3] = {0, 0, 0}, prevG[3] = {0, 0, 0}, prevA[3] = {0, 0, 0}, prevRawG[3] = {0, 0, 0}, prevRawA[3] = {0, 0, 0};
  double gyro[3] = {0, 0, 0}, accel[3] = {0, 0, 0};
  int steps = 0;
  enumIntegrationSteps(meas, model, timeStartUs, timeEndUs, [&](const float* gf, const float* af, double dtSec, bool newAccel, bool newGyro) {
    steps++;
    const double rawG[3] = {gf[0], gf[1], gf[2]}, rawA[3] = {af[0], af[1], af[2]};
    compensateAndJac(cmats, ix, rawG, rawA, gyro, accel, cmJ.data());
    double J96[54];
    const Rvp rvp = integrateJac(gyro, accel, dtSec, J96);
    for (int r = 0; r < 9; r++) for (int c = 0; c < 9; c++) rvp2Jac[r * n + c] = 0.0;
    gemmLd(J96, 6, cmJ.data(), nc, rvp2Jac.data() + 9, n, 9, 6, nc);
    if (newAccel && ix.gyroAccelDt != 0) {   // Q7
      double dG[3], dA[3];
      for (int i = 0; i < 3; i++) { dG[i] = gyro[i] - prevG[i]; dA[i] = accel[i] - prevA[i]; }
      if (newGyro) {
        double backG[3], backA[3], forwG[3], forwA[3];
        mat3vec(model.gyroScaleInv, rawG, forwG); mat3vec(model.accelScaleInv, prevRawA, forwA);
        mat3vec(model.gyroScaleInv, prevRawG, backG); mat3vec(model.accelScaleInv, rawA, backA);
        for (int i = 0; i < 3; i++) {
          forwG[i] -= model.gyroBias[i]; forwA[i] -= model.accelBias[i]; backG[i] -= model.gyroBias[i]; backA[i] -= model.accelBias[i];
          dG[i] = (backG[i] - prevG[i] + gyro[i] - forwG[i]) * 0.5;
          dA[i] = (backA[i] - prevA[i] + accel[i] - forwA[i]) * 0.5;
        }
      }
      double col[9];
      dRvpDLeft(rvp, dG, dA, col);
      for (int r = 0; r < 9; r++) rvp2Jac[r * n + 15 + ix.gyroAccelDt] = col[r];
    }
    for (int i = 0; i < 3; i++) { prevA[i] = accel[i]; prevG[i] = gyro[i]; prevRawA[i] = rawA[i]; prevRawG[i] = rawG[i]; }
    if (!have) {
      total = rvp; have = true;
      std::swap(rvpJac, rvp2Jac);
      for (int i = 0; i < 3; i++) { startG[i] = gyro[i]; startA[i] = accel[i]; }
    } else {
      std::swap(rvpJac, rvp1Jac);
      for (int r = 0; r < 9; r++) {
        for (int c = 0; c < 9; c++) rvp1Jac[r * n + c] = r == c ? 1.0 : 0.0;
        for (int c = 9; c < 15; c++) rvp1Jac[r * n + c] = 0.0;
      }
      total = combineJacs(total, rvp, rvp1Jac.data(), rvp2Jac.data(), rvpJac.data(), n);
    }
    double A[81], T[81];
    for (int r = 0; r < 9; r++) for (int c = 0; c < 9; c++) A[r * 9 + c] = rvpJac[r * n + c];
    gemmLd(A, 9, cov, 9, T, 9, 9, 9, 9);
    for (int r = 0; r < 9; r++) for (int c = 0; c < 9; c++) { double s = 0; for (int k = 0; k < 9; k++) s += T[r * 9 + k] * A[c * 9 + k]; cov[r * 9 + c] = s; }
    double t27[27];
    gemmLd(A, 9, fromG, 3, t27, 3, 9, 9, 3); std::memcpy(fromG, t27, sizeof t27);
    gemmLd(A, 9, fromA, 3, t27, 3, 9, 9, 3); std::memcpy(fromA, t27, sizeof t27);
    auto flush = [&](double* from, const double* var) {
      for (int r = 0; r < 9; r++) for (int c = 0; c < 9; c++) { double s = 0; for (int k = 0; k < 3; k++) s += from[r * 3 + k] * var[k] * from[c * 3 + k]; cov[r * 9 + c] += s; }
    };
    if (newGyro) { flush(fromG, gyroVar); for (double& x : fromG) x = 0.0; }
    if (newAccel) { flush(fromA, accelVar); for (double& x : fromA) x = 0.0; }
    for (int r = 0; r < 9; r++) for (int k = 0; k < 3; k++) { fromG[r * 3 + k] += rvpJac[r * n + 9 + k]; fromA[r * 3 + k] += rvpJac[r * n + 12 + k]; }
  });
  if (!have) throw std::runtime_error("computePreIntegration: empty interval");
  for (int r = 0; r < 9; r++) for (int c = 0; c < 9; c++) {
    double s = 0;
    for (int k = 0; k < 3; k++) s += fromG[r * 3 + k] * gyroVar[k] * fromG[c * 3 + k] + fromA[r * 3 + k] * accelVar[k] * fromA[c * 3 + k];
    cov[r * 9 + c] += s;
  }
  PreintResult out;
  std::memset(out.rec, 0, sizeof out.rec);
  for (int i = 0; i < 4; i++) out.rec[i] = total.q[i];
  for (int i = 0; i < 3; i++) { out.rec[4 + i] = total.dV[i]; out.rec[7 + i] = total.dP[i]; }
  out.rec[10] = total.dt;
  for (int r = 0; r < 9; r++) for (int c = 0; c < E; c++) out.rec[kPreintJ + r * 23 + c] = rvpJac[r * n + 15 + c];
  if (ix.refDt >= 0) {   // PreIntegration.cpp:260-265
    double nG[3] = {-startG[0], -startG[1], -startG[2]}, nA[3] = {-startA[0], -startA[1], -startA[2]}, s9[9], R[9], e1[3], e2[3];
    dRvpDLeft(total, nG, nA, s9);
    quatToMat(total.q, R);
    mat3vec(R, gyro, e1); mat3vec(R, accel, e2);
    for (int i = 0; i < 3; i++) {
      out.rec[kPreintJ + i * 23 + ix.refDt] = s9[i] + e1[i];
      out.rec[kPreintJ + (3 + i) * 23 + ix.refDt] = s9[3 + i] + e2[i];
      out.rec[kPreintJ + (6 + i) * 23 + ix.refDt] = s9[6 + i] + total.dV[i];
    }
  }
  for (int i = 0; i < 32; i++) out.rec[kPreintEval + i] = calib32[i];
  std::memcpy(out.cov, cov, sizeof cov);
  for (int i = 0; i < 3; i++) out.omegaAtEnd[i] = prevG[i];
  out.steps = steps;
  return out;
}

}  // namespace orc
