// Rolling-shutter table refresher: one thread per rig runs RollingShutterData::compute
// (dev_refresh.cuh) on the raw IMU-0 samples with the current calibration of the rig's window and writes
// the rig's table in place.  ~30 sequential integration steps per rig, a few thousand rigs: the kernel is
// latency-sized (tens of microseconds), its job is to keep the host and its 2 x 8 B x 632 x rigs of
// table traffic out of the iteration.
#include "dev_refresh.cuh"
#include "dev_preint.cuh"
#include "vb_internal.hpp"

namespace vb {

__global__ void __launch_bounds__(64) rsRefreshKernel(const int64_t* __restrict__ timeNs, const float* __restrict__ gyro,
                                                      const float* __restrict__ accel, int64_t nSamples,
                                                      const double* __restrict__ varData, const int64_t* __restrict__ calibOff,
                                                      const int64_t* __restrict__ midUs, int64_t halfUs, int64_t nRigs,
                                                      double* tables, int* status) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= nRigs) return;
  const int64_t off = calibOff[r];
  if (off < 0) return;  // rig without rolling-shutter observations: its table is left alone
  double calib[32];
#pragma unroll
  for (int i = 0; i < 32; i++) calib[i] = varData[off + i];
  const int st = rf::refreshRig(timeNs, gyro, accel, nSamples, calib, midUs[r], halfUs, tables + r * rf::kStride);
  if (st != rf::kOk) atomicMax(status, st);
}

void launchRsRefresh(const int64_t* timeNs, const float* gyro, const float* accel, int64_t nSamples, const double* varData,
                     const int64_t* calibOff, const int64_t* midUs, int64_t halfUs, int64_t nRigs, double* tables, int* status,
                     cudaStream_t st) {
  if (nRigs <= 0) return;
  const int block = 64;
  rsRefreshKernel<<<(unsigned)((nRigs + block - 1) / block), block, 0, st>>>(timeNs, gyro, accel, nSamples, varData, calibOff, midUs,
                                                                            halfUs, nRigs, tables, status);
}

// Preintegration refresher (second half): a group of 8 lanes per inertial factor (4 factors per warp)
// runs computePreIntegration over the raw samples of the factor's IMU between its two rigs, with the CURRENT calibration
// of the factor's window, and writes the record the inertial kernels read.  The scan over the ~100-250 integration
// pieces is sequential; inside a piece the work is spread over the group's lanes: columns of the 9 x (15 + E) Jacobian
// chain (combineJacs is independent per column; the step Jacobian's column is produced on the fly), entries of the
// 9 x 9 covariance and of the two 9 x 3 pending-noise maps, all held in shared memory.  The scalar part of a piece
// (compensation, integrate + its 9 x 6 Jacobian, the combine context) is computed redundantly by the 8 lanes.
// Arithmetic per column / entry is that of pi::preintegrate (dev_preint.cuh), the sequential form the CPU test compiles.
// Measured at C3 (12 k factors, 200 pieces each): one thread per factor 23 ms; see profiles/ for this kernel.
constexpr int kPreGW = 8;                 // lanes per factor
constexpr int kPreWarps = 2;              // warps per CTA
constexpr int kPreGroups = kPreWarps * 32 / kPreGW;
constexpr int kPreSm = 9 * pi::kMaxN + 81 * 3 + 54;   // J, A, cov, T, [FG | FA]
__global__ void __launch_bounds__(32 * kPreWarps) preintRefreshKernel(const int64_t* __restrict__ timeNs, const float* __restrict__ gyro,
                                                          const float* __restrict__ accel, int64_t nS, const double* __restrict__ varData,
                                                          const int64_t* __restrict__ calibOff, const uint32_t* __restrict__ calibFlags,
                                                          const int64_t* __restrict__ startUs, const int64_t* __restrict__ endUs, int64_t nFactors,
                                                          const double* __restrict__ noise6, double* records, double* covOut,
                                                          double* omegaOut, int* status) {
  __shared__ double sh[kPreGroups][kPreSm];
  const int lane = threadIdx.x & 31, gl = lane & (kPreGW - 1), grp = threadIdx.x / kPreGW;
  const unsigned gmask = ((1u << kPreGW) - 1u) << (lane & ~(kPreGW - 1));
  const int64_t k = blockIdx.x * (int64_t)kPreGroups + grp;
  if (k >= nFactors) return;
  const int64_t off = calibOff[k];
  if (off < 0) return;
  double* J = sh[grp];                       // [9][n]
  double* A = J + 9 * pi::kMaxN; double* cov = A + 81; double* T = cov + 81; double* FG = T + 81; double* FA = FG + 27;
  double calib[32];
#pragma unroll
  for (int i = 0; i < 32; i++) calib[i] = varData[off + i];
  const double gv[3] = {noise6[0], noise6[1], noise6[2]}, av[3] = {noise6[3], noise6[4], noise6[5]};
  pi::Idx ix;
  pi::makeIdx(calibFlags[k], ix);
  const int E = ix.size, n = 15 + E;
  rf::ImuModel mdl;
  rf::imuModel(calib, mdl);
  pi::CalibM cm;
  pi::calibMats(calib, cm);
  const int64_t t0 = startUs[k] * 1000, t1 = endUs[k] * 1000;
  const int64_t gFirst = rf::firstAfter(timeNs, nS, t0 + mdl.dtGyroNs + rf::kMarginNs), gLast = rf::firstAfter(timeNs, nS, t1 + mdl.dtGyroNs - rf::kMarginNs);
  const int64_t aFirst = rf::firstAfter(timeNs, nS, t0 + mdl.dtAccelNs + rf::kMarginNs), aLast = rf::firstAfter(timeNs, nS, t1 + mdl.dtAccelNs - rf::kMarginNs);
  if (gFirst <= 0 || aFirst <= 0 || gLast >= nS || aLast >= nS) { if (gl == 0) atomicMax(status, pi::kNoMargin); return; }
  for (int e = gl; e < 81; e += kPreGW) cov[e] = 0.0;
  for (int e = gl; e < 54; e += kPreGW) FG[e] = 0.0;
  for (int e = gl; e < 9 * n; e += kPreGW) J[e] = 0.0;
  __syncwarp(gmask);
  rf::Seg total;
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
    pi::mv3(cm.gS, rawG, sG);
    pi::mv3(cm.aS, rawA, sA);
    for (int i = 0; i < 3; i++) { g[i] = sG[i] - cm.gb[i]; a[i] = sA[i] - cm.ab[i]; }
    rf::Seg piece;
    double J96[54];
    pi::integrateJ(g, a, dt, piece, J96);
    const bool q7 = newAccel && ix.gadt != 0;   // Q7 of oracle/preint.hpp
    double q7col[9];
    if (q7) {
      double dG[3], dA[3];
      for (int i = 0; i < 3; i++) { dG[i] = g[i] - prevG[i]; dA[i] = a[i] - prevA[i]; }
      if (newGyro) {
        double fG[3], fA[3], bG[3], bA[3];
        for (int r = 0; r < 3; r++) {
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
      pi::dLeft(piece, dG, dA, q7col);
    }
    for (int i = 0; i < 3; i++) { prevA[i] = a[i]; prevG[i] = g[i]; prevRawA[i] = rawA[i]; prevRawG[i] = rawG[i]; }
    double aR[9], aRbV[3], aRbP[3];
    pi::quatToR(total.q, aR);
    pi::mv3(aR, piece.dV, aRbV);
    pi::mv3(aR, piece.dP, aRbP);
    for (int col = gl; col < n; col += kPreGW) {   // this lane's columns of the chain
      double b9[9];
      if (col < 9) { for (int r = 0; r < 9; r++) b9[r] = 0.0; }
      else {
        double c6[6];
        pi::compCol(cm, ix, rawG, rawA, sG, sA, col - 9, c6);
        for (int r = 0; r < 9; r++) {
          double s = 0.0;
          for (int kk = 0; kk < 6; kk++) s += J96[r * 6 + kk] * c6[kk];
          b9[r] = s;
        }
      }
      if (q7 && col == 15 + ix.gadt) for (int r = 0; r < 9; r++) b9[r] = q7col[r];
      if (!have) { for (int r = 0; r < 9; r++) J[r * n + col] = b9[r]; continue; }
      double at[3], avv[3], ap[3];
      for (int i = 0; i < 3; i++) {
        if (col < 9) { at[i] = (col == i) ? 1.0 : 0.0; avv[i] = (col == 3 + i) ? 1.0 : 0.0; ap[i] = (col == 6 + i) ? 1.0 : 0.0; }
        else if (col < 15) { at[i] = avv[i] = ap[i] = 0.0; }
        else { at[i] = J[i * n + col]; avv[i] = J[(3 + i) * n + col]; ap[i] = J[(6 + i) * n + col]; }
      }
      double Rt[3], Rv[3], Rp[3], hv[3], hp[3];
      pi::mv3(aR, b9, Rt); pi::mv3(aR, b9 + 3, Rv); pi::mv3(aR, b9 + 6, Rp);
      rf::cross(at, aRbV, hv);
      rf::cross(at, aRbP, hp);
      for (int i = 0; i < 3; i++) {
        J[i * n + col] = at[i] + Rt[i];
        J[(3 + i) * n + col] = avv[i] + hv[i] + Rv[i];
        J[(6 + i) * n + col] = ap[i] + avv[i] * piece.dt + hp[i] + Rp[i];
      }
    }
    if (!have) {
      total = piece; have = true;
      for (int i = 0; i < 3; i++) { startG[i] = g[i]; startA[i] = a[i]; }
    } else {
      rf::combine(total, piece, total);
    }
    __syncwarp(gmask);
    // covariance and pending noise maps, entry-parallel (PreIntegration.cpp:231-250)
    for (int e = gl; e < 81; e += kPreGW) { const int r = e / 9, c = e % 9; A[e] = J[r * n + c]; }
    __syncwarp(gmask);
    for (int e = gl; e < 81; e += kPreGW) { const int r = e / 9, c = e % 9; double s = 0.0; for (int kk = 0; kk < 9; kk++) s += A[r * 9 + kk] * cov[kk * 9 + c]; T[e] = s; }
    double fn[7];   // new [FG | FA] entries of this lane (54 / 8 -> 7)
    {
      int q = 0;
      for (int e = gl; e < 54; e += kPreGW, q++) {
        const double* F = e < 27 ? FG : FA; const int ee = e < 27 ? e : e - 27; const int r = ee / 3, c = ee % 3;
        double s = 0.0;
        for (int kk = 0; kk < 9; kk++) s += A[r * 9 + kk] * F[kk * 3 + c];
        fn[q] = s;
      }
    }
    __syncwarp(gmask);
    { int q = 0; for (int e = gl; e < 54; e += kPreGW, q++) FG[e] = fn[q]; }   // FG and FA are contiguous
    for (int e = gl; e < 81; e += kPreGW) { const int r = e / 9, c = e % 9; double s = 0.0; for (int kk = 0; kk < 9; kk++) s += T[r * 9 + kk] * A[c * 9 + kk]; cov[e] = s; }
    __syncwarp(gmask);
    if (newGyro) {
      for (int e = gl; e < 81; e += kPreGW) { const int r = e / 9, c = e % 9; double s = 0.0; for (int kk = 0; kk < 3; kk++) s += FG[r * 3 + kk] * gv[kk] * FG[c * 3 + kk]; cov[e] += s; }
      __syncwarp(gmask);
      for (int e = gl; e < 27; e += kPreGW) FG[e] = 0.0;
    }
    if (newAccel) {
      for (int e = gl; e < 81; e += kPreGW) { const int r = e / 9, c = e % 9; double s = 0.0; for (int kk = 0; kk < 3; kk++) s += FA[r * 3 + kk] * av[kk] * FA[c * 3 + kk]; cov[e] += s; }
      __syncwarp(gmask);
      for (int e = gl; e < 27; e += kPreGW) FA[e] = 0.0;
    }
    __syncwarp(gmask);
    for (int e = gl; e < 27; e += kPreGW) { const int r = e / 3, c = e % 3; FG[e] += J[r * n + 9 + c]; FA[e] += J[r * n + 12 + c]; }
    __syncwarp(gmask);
  }
  if (!have) { if (gl == 0) atomicMax(status, pi::kEmpty); return; }
  for (int e = gl; e < 81; e += kPreGW) {
    const int r = e / 9, c = e % 9;
    double s = 0.0;
    for (int kk = 0; kk < 3; kk++) s += FG[r * 3 + kk] * gv[kk] * FG[c * 3 + kk] + FA[r * 3 + kk] * av[kk] * FA[c * 3 + kk];
    covOut[k * 81 + e] = cov[e] + s;
  }
  double* rec = records + k * pi::kRec;
  if (gl < 4) rec[gl] = total.q[gl];
  if (gl < 3) { rec[4 + gl] = total.dV[gl]; rec[7 + gl] = total.dP[gl]; omegaOut[k * 3 + gl] = prevG[gl]; }
  if (gl == 0) rec[10] = total.dt;
  for (int i = gl; i < 32; i += kPreGW) rec[pi::kRecEval + i] = calib[i];
  // calibration columns: column c of the record = chain column 15 + c; the reference-IMU time-offset column is
  // replaced by dRvp_dStartTime + dRvp_dEndTime (PreIntegration.cpp:260-265)
  double s9[9], Rm[9], e1[3], e2[3];
  {
    const double nG[3] = {-startG[0], -startG[1], -startG[2]}, nA[3] = {-startA[0], -startA[1], -startA[2]};
    pi::dLeft(total, nG, nA, s9);
    pi::quatToR(total.q, Rm);
    pi::mv3(Rm, g, e1); pi::mv3(Rm, a, e2);
  }
  for (int c = gl; c < pi::kMaxE; c += kPreGW)
    for (int r = 0; r < 9; r++) {
      double v = c < E ? J[r * n + 15 + c] : 0.0;
      if (c == ix.rdt) v = r < 3 ? s9[r] + e1[r] : r < 6 ? s9[r] + e2[r - 3] : s9[r] + total.dV[r - 6];
      rec[pi::kRecJ + r * pi::kMaxE + c] = v;
    }
}

void launchPreintRefresh(const int64_t* timeNs, const float* gyro, const float* accel, int64_t nSamples, const double* varData,
                         const int64_t* calibOff, const uint32_t* calibFlags, const int64_t* startUs, const int64_t* endUs, int64_t nFactors,
                         const double* noise6, double* scratch, double* records, double* cov, double* omegaAtEnd, int* status, cudaStream_t st) {
  if (nFactors <= 0) return;
  (void)scratch;
  preintRefreshKernel<<<(unsigned)((nFactors + kPreGroups - 1) / kPreGroups), 32 * kPreWarps, 0, st>>>(timeNs, gyro, accel, nSamples, varData, calibOff,
                                                                                                   calibFlags, startUs, endUs, nFactors, noise6, records,
                                                                                                   cov, omegaAtEnd, status);
}

}  // namespace vb
