// Per-iteration refresher of the preintegration records on the device (second half):
// computePreIntegration (lib/motion/preintegration/PreIntegration.cpp:136-275) with
// getCompensatedImuMeasurementAndJac (CompensateJac.cpp:146-249), integrate(gyro, accel, dt, &J) and combineJacs
// (MotionIntegral.cpp:52-75,162-226), reached every LM iteration under --recompute-preint through
// SingleSessionAdapter::regenerateAllPreintegrationsFromImuMeasurements (viba/single_session/InertialFactors.cpp:19-70).
//
// One thread refreshes one inertial factor: a sequential scan over the ~100-250 integration pieces between two rigs,
// carrying the 9 x (15 + E) Jacobian chain (E <= 23 calibration error states), the 9 x 9 covariance and the two 9 x 3
// pending-noise maps.  The step Jacobian rvp2Jac is never materialised: its columns are produced on the fly inside the
// column-wise combine (combineJacs is independent per column).  The record is written straight into the factor store's
// aux table (kPreintStride doubles: R dV dP dt J(9x23) evalPoint(32)), which the inertial kernels read.
//
// Like dev_refresh.cuh the file is free of device-only constructs, so tests/cpp/preint_host.cpp compiles it as host
// code and compares it with the oracle without a GPU; the library only ever runs it inside preintRefreshKernel.
#pragma once
#include "dev_refresh.cuh"

namespace vb {
namespace pi {

using rf::Seg;

constexpr int kRec = 250, kRecJ = 11, kRecEval = 218, kMaxE = 23, kMaxN = 15 + kMaxE;

struct Idx {  // lib/motion/imu_types/ImuCalibrationJacobianIndices.h:37-96
  int gb, ab, gs, as, gn, an, rdt, gadt, size;
};
VB_RF void makeIdx(uint32_t f, Idx& ix) {
  int r = 0;
  ix.gb = (f & 1) ? r : -1; r += (f & 1) ? 3 : 0;
  ix.ab = (f & 2) ? r : -1; r += (f & 2) ? 3 : 0;
  ix.gs = (f & 4) ? r : -1; r += (f & 4) ? 3 : 0;
  ix.as = (f & 8) ? r : -1; r += (f & 8) ? 3 : 0;
  ix.gn = (f & 16) ? r : -1; r += (f & 16) ? 6 : 0;
  ix.an = (f & 32) ? r : -1; r += (f & 32) ? 3 : 0;
  ix.rdt = (f & 64) ? r : -1; r += (f & 64) ? 1 : 0;
  ix.gadt = (f & 128) ? r : -1; r += (f & 128) ? 1 : 0;
  ix.size = r;
}

struct CalibM {  // CompensateJac.cpp:160-166; matrices row-major
  double gNinv[9], aNinv[9], gS[9], aS[9], gN[9], aN[9], gb[3], ab[3];
};
VB_RF void calibMats(const double* c, CalibM& m) {
  for (int r = 0; r < 3; r++)
    for (int k = 0; k < 3; k++) { m.gN[r * 3 + k] = c[12 + k * 3 + r]; m.aN[r * 3 + k] = c[21 + k * 3 + r]; }
  rf::inv3(m.gN, m.gNinv);
  rf::inv3(m.aN, m.aNinv);
  for (int r = 0; r < 3; r++)
    for (int k = 0; k < 3; k++) { m.gS[r * 3 + k] = m.gNinv[r * 3 + k] / c[k]; m.aS[r * 3 + k] = m.aNinv[r * 3 + k] / c[3 + k]; }
  for (int i = 0; i < 3; i++) { m.gb[i] = c[6 + i]; m.ab[i] = c[9 + i]; }
}
VB_RF void mv3(const double* M, const double* v, double* o) {
  for (int r = 0; r < 3; r++) o[r] = M[r * 3] * v[0] + M[r * 3 + 1] * v[1] + M[r * 3 + 2] * v[2];
}

// column k of the 6 x (6 + E) matrix [measJac | calibJac] of the compensation (CompensateJac.cpp:168-248);
// sG / sA = scale matrices applied to the raw gyro / accel samples
VB_RF void compCol(const CalibM& m, const Idx& ix, const double* rawG, const double* rawA, const double* sG, const double* sA, int k, double* o) {
  for (int i = 0; i < 6; i++) o[i] = 0.0;
  if (k < 3) { for (int r = 0; r < 3; r++) o[r] = m.gS[r * 3 + k]; return; }
  if (k < 6) { for (int r = 0; r < 3; r++) o[3 + r] = m.aS[r * 3 + k - 3]; return; }
  const int e = k - 6;
  if (ix.gb >= 0 && e >= ix.gb && e < ix.gb + 3) { o[e - ix.gb] = -1.0; return; }
  if (ix.ab >= 0 && e >= ix.ab && e < ix.ab + 3) { o[3 + e - ix.ab] = -1.0; return; }
  if (ix.gs >= 0 && e >= ix.gs && e < ix.gs + 3) { const int j = e - ix.gs; for (int r = 0; r < 3; r++) o[r] = m.gNinv[r * 3 + j] * rawG[j]; return; }
  if (ix.as >= 0 && e >= ix.as && e < ix.as + 3) { const int j = e - ix.as; for (int r = 0; r < 3; r++) o[3 + r] = m.aNinv[r * 3 + j] * rawA[j]; return; }
  if (ix.gn >= 0 && e >= ix.gn && e < ix.gn + 6) {
    const int i = e - ix.gn;
    const int r = i >> 1, c = (i == 0) ? 1 : (i == 1) ? 2 : (i == 2) ? 0 : (i == 3) ? 2 : (i == 4) ? 0 : 1;  // rows 0,0,1,1,2,2 / cols 1,2,0,2,0,1
    const double f = sG[r] * (-m.gN[r * 3 + c] / m.gN[r * 3 + r]) + sG[c];
    for (int q = 0; q < 3; q++) o[q] = -m.gNinv[q * 3 + r] * f;
    return;
  }
  if (ix.an >= 0 && e >= ix.an && e < ix.an + 3) {
    const int i = e - ix.an;
    const int r = (i == 2) ? 1 : 0, c = (i == 0) ? 1 : 2;   // rows 0,0,1 / cols 1,2,2
    const double f = sA[r] * (-m.aN[r * 3 + c] / m.aN[r * 3 + r]) + sA[c];
    for (int q = 0; q < 3; q++) o[3 + q] = -m.aNinv[q * 3 + r] * f;
    return;
  }
  // time-offset columns: zero here (filled by the caller)
}

VB_RF void hatm(const double* v, double* H) {
  H[0] = 0; H[1] = -v[2]; H[2] = v[1]; H[3] = v[2]; H[4] = 0; H[5] = -v[0]; H[6] = -v[1]; H[7] = v[0]; H[8] = 0;
}
VB_RF void mm3(const double* A, const double* B, double* C) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) C[i * 3 + j] = A[i * 3] * B[j] + A[i * 3 + 1] * B[3 + j] + A[i * 3 + 2] * B[6 + j];
}
VB_RF void quatToR(const double* q, double* R) {
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  R[0] = 1 - 2 * (y * y + z * z); R[1] = 2 * (x * y - z * w); R[2] = 2 * (x * z + y * w);
  R[3] = 2 * (x * y + z * w); R[4] = 1 - 2 * (x * x + z * z); R[5] = 2 * (y * z - x * w);
  R[6] = 2 * (x * z - y * w); R[7] = 2 * (y * z + x * w); R[8] = 1 - 2 * (x * x + y * y);
}

// MotionIntegral.cpp:162-226: the segment and its 9 x 6 Jacobian wrt (gyro, accel), row-major
VB_RF void integrateJ(const double* gyro, const double* accel, double dt, Seg& r, double* J) {
  double om[3], up[3], upd[3];
  for (int i = 0; i < 3; i++) { om[i] = gyro[i] * dt; up[i] = accel[i] * dt; upd[i] = up[i] * dt; }
  rf::expQ(om, r.q);
  const double th2 = om[0] * om[0] + om[1] * om[1] + om[2] * om[2], th = sqrt(th2), th4 = th2 * th2;
  double c1, c2, c3, d1, d2, d3;
  if (th < 1e-3) {
    c1 = 0.5 - th2 / 24.0 + th4 / 729.0;   // F6 = 729 (sic), MotionIntegral.cpp:82
    c2 = 1.0 / 6.0 - th2 / 120.0 + th4 / 5040.0;
    c3 = 1.0 / 24.0 - th2 / 729.0 + th4 / 40320.0;
    d1 = -(2.0 / 24.0) + th2 * (4.0 / 729.0) + th4 * (6.0 / 40320.0);
    d2 = -(2.0 / 120.0) + th2 * (4.0 / 5040.0) + th4 * (6.0 / 362880.0);
    d3 = -(2.0 / 729.0) + th2 * (4.0 / 40320.0) + th4 * (6.0 / 3628800.0);
  } else {
    const double s = sin(th) / th, m = (1.0 - cos(th)) / th2;
    c1 = m; c2 = (1.0 - s) / th2; c3 = (0.5 - m) / th2;
    d1 = (s - 2.0 * m) / th2; d2 = (m - 3.0 * c2) / th2; d3 = (-1.0 - s + 4.0 * m) / th4;
  }
  double W[9], W2[9], U2V[9], U2P[9];
  hatm(om, W);
  mm3(W, W, W2);
  for (int i = 0; i < 9; i++) { U2V[i] = c1 * W[i] + c2 * W2[i]; U2P[i] = c2 * W[i] + c3 * W2[i]; }
  U2V[0] += 1; U2V[4] += 1; U2V[8] += 1;
  U2P[0] += 0.5; U2P[4] += 0.5; U2P[8] += 0.5;
  mv3(U2V, up, r.dV);
  mv3(U2P, upd, r.dP);
  r.dt = dt;
  double nup[3] = {-up[0], -up[1], -up[2]}, DwXu[9], oxu[3], H2[9], WD[9], DwXwXu[9], M1[9], M2[9], VD1[3], PD1[3];
  hatm(nup, DwXu);
  for (int i = 0; i < 9; i++) DwXu[i] *= dt;
  rf::cross(om, up, oxu);
  const double noxu[3] = {-oxu[0], -oxu[1], -oxu[2]};
  hatm(noxu, H2);
  mm3(W, DwXu, WD);
  for (int i = 0; i < 9; i++) { DwXwXu[i] = H2[i] * dt + WD[i]; M1[i] = d1 * W[i] + d2 * W2[i]; M2[i] = d2 * W[i] + d3 * W2[i]; }
  mv3(M1, up, VD1);
  mv3(M2, upd, PD1);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      J[i * 6 + j] = U2V[i * 3 + j] * dt;
      J[i * 6 + 3 + j] = 0.0;
      J[(3 + i) * 6 + j] = VD1[i] * om[j] * dt + c1 * DwXu[i * 3 + j] + c2 * DwXwXu[i * 3 + j];
      J[(6 + i) * 6 + j] = PD1[i] * om[j] * dt + (c2 * DwXu[i * 3 + j] + c3 * DwXwXu[i * 3 + j]) * dt;
      J[(3 + i) * 6 + 3 + j] = U2V[i * 3 + j] * dt;
      J[(6 + i) * 6 + 3 + j] = U2P[i * 3 + j] * dt * dt;
    }
}

// PreIntegration.cpp:113-120
VB_RF void dLeft(const Seg& s, const double* g, const double* a, double* o) {
  double t[3];
  for (int i = 0; i < 3; i++) o[i] = g[i];
  rf::cross(g, s.dV, t);   // hat(-dV) g = g x dV
  for (int i = 0; i < 3; i++) o[3 + i] = t[i] + a[i];
  rf::cross(g, s.dP, t);
  for (int i = 0; i < 3; i++) o[6 + i] = a[i] * s.dt + t[i];
}

enum { kOk = 0, kNoMargin = 1, kEmpty = 4 };

// computePreIntegration for one factor.  J (9 x kMaxN, row-major, leading dimension n = 15 + E) is caller-provided
// scratch.  Outputs: rec[250], cov[81] (may be null), omegaAtEnd[3] (may be null).
VB_RF int preintegrate(const int64_t* timeNs, const float* gyro, const float* accel, int64_t nS, const double* calib32, uint32_t flags,
                       const double* gyroVar, const double* accelVar, int64_t startUs, int64_t endUs, double* J, double* rec, double* covOut,
                       double* omegaOut) {
  Idx ix;
  makeIdx(flags, ix);
  const int E = ix.size, n = 15 + E;
  rf::ImuModel mdl;
  rf::imuModel(calib32, mdl);
  CalibM cm;
  calibMats(calib32, cm);
  const int64_t t0 = startUs * 1000, t1 = endUs * 1000;
  const int64_t gFirst = rf::firstAfter(timeNs, nS, t0 + mdl.dtGyroNs + rf::kMarginNs), gLast = rf::firstAfter(timeNs, nS, t1 + mdl.dtGyroNs - rf::kMarginNs);
  const int64_t aFirst = rf::firstAfter(timeNs, nS, t0 + mdl.dtAccelNs + rf::kMarginNs), aLast = rf::firstAfter(timeNs, nS, t1 + mdl.dtAccelNs - rf::kMarginNs);
  if (gFirst <= 0 || aFirst <= 0 || gLast >= nS || aLast >= nS) return kNoMargin;
  double cov[81], FG[27], FA[27];
  for (int i = 0; i < 81; i++) cov[i] = 0.0;
  for (int i = 0; i < 27; i++) FG[i] = FA[i] = 0.0;
  Seg total;
  total.q[0] = total.q[1] = total.q[2] = 0.0; total.q[3] = 1.0;
  for (int i = 0; i < 3; i++) total.dV[i] = total.dP[i] = 0.0;
  total.dt = 0.0;
  bool have = false;
  double startG[3] = {0, 0, 0}, startA[3] = {0, 0, 0}, prevG[3] = {0, 0, 0}, prevA[3] = {0, 0, 0}, prevRawG[3] = {0, 0, 0}, prevRawA[3] = {0, 0, 0};
  double g[3] = {0, 0, 0}, a[3] = {0, 0, 0};
  int64_t prevStamp = t0, gi = gFirst, ai = aFirst;
  while (gi <= gLast && ai <= aLast) {
    const int64_t gEnd = timeNs[gi] - mdl.dtGyroNs, aEnd = timeNs[ai] - mdl.dtAccelNs;
    const int64_t pieceEnd = gEnd < aEnd ? gEnd : aEnd;
    const bool started = gi > gFirst || ai > aFirst;
    const bool newAccel = started && (timeNs[ai - 1] - mdl.dtAccelNs == prevStamp);
    const bool newGyro = started && (timeNs[gi - 1] - mdl.dtGyroNs == prevStamp);
    const bool last = gi >= gLast && ai >= aLast;
    const int64_t stamp = last ? t1 : pieceEnd;
    const double dt = (double)(stamp - prevStamp) * 1e-9;
    prevStamp = stamp;
    const double rawG[3] = {(double)gyro[gi * 3], (double)gyro[gi * 3 + 1], (double)gyro[gi * 3 + 2]};
    const double rawA[3] = {(double)accel[ai * 3], (double)accel[ai * 3 + 1], (double)accel[ai * 3 + 2]};
    gi += (gEnd == pieceEnd);
    ai += (aEnd == pieceEnd);
    double sG[3], sA[3];
    mv3(cm.gS, rawG, sG);
    mv3(cm.aS, rawA, sA);
    for (int i = 0; i < 3; i++) { g[i] = sG[i] - cm.gb[i]; a[i] = sA[i] - cm.ab[i]; }
    Seg piece;
    double J96[54];
    integrateJ(g, a, dt, piece, J96);
    // Q7: `jacInd.gyroAccelTimeOffsetIdx()` is tested as a boolean in the reference (PreIntegration.cpp:197)
    const bool q7 = newAccel && ix.gadt != 0;
    double q7col[9];
    if (q7) {
      double dG[3], dA[3];
      for (int i = 0; i < 3; i++) { dG[i] = g[i] - prevG[i]; dA[i] = a[i] - prevA[i]; }
      if (newGyro) {
        double fG[3], fA[3], bG[3], bA[3];
        for (int r = 0; r < 3; r++) {   // getCompensatedImuMeasurement: (diag(scale) nonorth)^-1 raw - bias
          fG[r] = mdl.gyroInv[r * 3] * rawG[0] + mdl.gyroInv[r * 3 + 1] * rawG[1] + mdl.gyroInv[r * 3 + 2] * rawG[2] - mdl.gyroBias[r];
          fA[r] = mdl.accelInv[r * 3] * prevRawA[0] + mdl.accelInv[r * 3 + 1] * prevRawA[1] + mdl.accelInv[r * 3 + 2] * prevRawA[2] - mdl.accelBias[r];
          bG[r] = mdl.gyroInv[r * 3] * prevRawG[0] + mdl.gyroInv[r * 3 + 1] * prevRawG[1] + mdl.gyroInv[r * 3 + 2] * prevRawG[2] - mdl.gyroBias[r];
          bA[r] = mdl.accelInv[r * 3] * rawA[0] + mdl.accelInv[r * 3 + 1] * rawA[1] + mdl.accelInv[r * 3 + 2] * rawA[2] - mdl.accelBias[r];
        }
        for (int i = 0; i < 3; i++) {
          dG[i] = (bG[i] - prevG[i] + g[i] - fG[i]) * 0.5;
          dA[i] = (bA[i] - prevA[i] + a[i] - fA[i]) * 0.5;
        }
      }
      dLeft(piece, dG, dA, q7col);
    }
    for (int i = 0; i < 3; i++) { prevA[i] = a[i]; prevG[i] = g[i]; prevRawA[i] = rawA[i]; prevRawG[i] = rawG[i]; }
    // combine context (MotionIntegral.cpp:58-64), taken from `total` BEFORE it is advanced
    double aR[9], aRbV[3], aRbP[3];
    quatToR(total.q, aR);
    mv3(aR, piece.dV, aRbV);
    mv3(aR, piece.dP, aRbP);
    for (int col = 0; col < n; col++) {
      double b9[9];
      if (col < 9) { for (int r = 0; r < 9; r++) b9[r] = 0.0; }
      else {
        double c6[6];
        compCol(cm, ix, rawG, rawA, sG, sA, col - 9, c6);
        for (int r = 0; r < 9; r++) {
          double s = 0.0;
          for (int k = 0; k < 6; k++) s += J96[r * 6 + k] * c6[k];
          b9[r] = s;
        }
      }
      if (q7 && col == 15 + ix.gadt) for (int r = 0; r < 9; r++) b9[r] = q7col[r];
      if (!have) { for (int r = 0; r < 9; r++) J[r * n + col] = b9[r]; continue; }
      double at[3], av[3], ap[3];
      for (int i = 0; i < 3; i++) {
        if (col < 9) { at[i] = (col == i) ? 1.0 : 0.0; av[i] = (col == 3 + i) ? 1.0 : 0.0; ap[i] = (col == 6 + i) ? 1.0 : 0.0; }
        else if (col < 15) { at[i] = av[i] = ap[i] = 0.0; }
        else { at[i] = J[i * n + col]; av[i] = J[(3 + i) * n + col]; ap[i] = J[(6 + i) * n + col]; }
      }
      double Rt[3], Rv[3], Rp[3], hv[3], hp[3];
      mv3(aR, b9, Rt); mv3(aR, b9 + 3, Rv); mv3(aR, b9 + 6, Rp);
      rf::cross(at, aRbV, hv);   // hat(-aRbV) at = at x aRbV
      rf::cross(at, aRbP, hp);
      for (int i = 0; i < 3; i++) {
        J[i * n + col] = at[i] + Rt[i];
        J[(3 + i) * n + col] = av[i] + hv[i] + Rv[i];
        J[(6 + i) * n + col] = ap[i] + av[i] * piece.dt + hp[i] + Rp[i];
      }
    }
    if (!have) {
      total = piece; have = true;
      for (int i = 0; i < 3; i++) { startG[i] = g[i]; startA[i] = a[i]; }
    } else {
      rf::combine(total, piece, total);
    }
    // covariance and pending noise maps (PreIntegration.cpp:231-250)
    double A[81], T[81];
    for (int r = 0; r < 9; r++) for (int c = 0; c < 9; c++) A[r * 9 + c] = J[r * n + c];
    for (int r = 0; r < 9; r++) for (int c = 0; c < 9; c++) { double s = 0.0; for (int k = 0; k < 9; k++) s += A[r * 9 + k] * cov[k * 9 + c]; T[r * 9 + c] = s; }
    for (int r = 0; r < 9; r++) for (int c = 0; c < 9; c++) { double s = 0.0; for (int k = 0; k < 9; k++) s += T[r * 9 + k] * A[c * 9 + k]; cov[r * 9 + c] = s; }
    for (int r = 0; r < 9; r++) for (int c = 0; c < 3; c++) { double s = 0.0, u = 0.0; for (int k = 0; k < 9; k++) { s += A[r * 9 + k] * FG[k * 3 + c]; u += A[r * 9 + k] * FA[k * 3 + c]; } T[r * 3 + c] = s; T[27 + r * 3 + c] = u; }
    for (int i = 0; i < 27; i++) { FG[i] = T[i]; FA[i] = T[27 + i]; }
    if (newGyro) {
      for (int r = 0; r < 9; r++) for (int c = 0; c < 9; c++) { double s = 0.0; for (int k = 0; k < 3; k++) s += FG[r * 3 + k] * gyroVar[k] * FG[c * 3 + k]; cov[r * 9 + c] += s; }
      for (int i = 0; i < 27; i++) FG[i] = 0.0;
    }
    if (newAccel) {
      for (int r = 0; r < 9; r++) for (int c = 0; c < 9; c++) { double s = 0.0; for (int k = 0; k < 3; k++) s += FA[r * 3 + k] * accelVar[k] * FA[c * 3 + k]; cov[r * 9 + c] += s; }
      for (int i = 0; i < 27; i++) FA[i] = 0.0;
    }
    for (int r = 0; r < 9; r++) for (int k = 0; k < 3; k++) { FG[r * 3 + k] += J[r * n + 9 + k]; FA[r * 3 + k] += J[r * n + 12 + k]; }
  }
  if (!have) return kEmpty;
  for (int r = 0; r < 9; r++) for (int c = 0; c < 9; c++) {
    double s = 0.0;
    for (int k = 0; k < 3; k++) s += FG[r * 3 + k] * gyroVar[k] * FG[c * 3 + k] + FA[r * 3 + k] * accelVar[k] * FA[c * 3 + k];
    cov[r * 9 + c] += s;
  }
  for (int i = 0; i < 4; i++) rec[i] = total.q[i];
  for (int i = 0; i < 3; i++) { rec[4 + i] = total.dV[i]; rec[7 + i] = total.dP[i]; }
  rec[10] = total.dt;
  for (int r = 0; r < 9; r++) for (int c = 0; c < kMaxE; c++) rec[kRecJ + r * kMaxE + c] = c < E ? J[r * n + 15 + c] : 0.0;
  if (ix.rdt >= 0) {   // PreIntegration.cpp:260-265
    const double nG[3] = {-startG[0], -startG[1], -startG[2]}, nA[3] = {-startA[0], -startA[1], -startA[2]};
    double s9[9], R[9], e1[3], e2[3];
    dLeft(total, nG, nA, s9);
    quatToR(total.q, R);
    mv3(R, g, e1); mv3(R, a, e2);
    for (int i = 0; i < 3; i++) {
      rec[kRecJ + i * kMaxE + ix.rdt] = s9[i] + e1[i];
      rec[kRecJ + (3 + i) * kMaxE + ix.rdt] = s9[3 + i] + e2[i];
      rec[kRecJ + (6 + i) * kMaxE + ix.rdt] = s9[6 + i] + total.dV[i];
    }
  }
  for (int i = 0; i < 32; i++) rec[kRecEval + i] = calib32[i];
  if (covOut) for (int i = 0; i < 81; i++) covOut[i] = cov[i];
  if (omegaOut) for (int i = 0; i < 3; i++) omegaOut[i] = prevG[i];
  return kOk;
}

}  // namespace pi
}  // namespace vb
