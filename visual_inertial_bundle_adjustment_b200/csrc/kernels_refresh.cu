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

// Preintegration refresher (second half): one thread per inertial factor runs computePreIntegration
// (dev_preint.cuh) over the raw samples of the factor's IMU between its two rigs, with the CURRENT calibration of the
// factor's window, and writes the record the inertial kernels read.  The 9 x (15 + E) Jacobian chain lives in a
// per-thread slab of global scratch (L1/L2 resident: 2.7 KB per thread), everything else in registers / local memory.
__global__ void __launch_bounds__(64) preintRefreshKernel(const int64_t* __restrict__ timeNs, const float* __restrict__ gyro,
                                                          const float* __restrict__ accel, int64_t nSamples, const double* __restrict__ varData,
                                                          const int64_t* __restrict__ calibOff, const uint32_t* __restrict__ calibFlags,
                                                          const int64_t* __restrict__ startUs, const int64_t* __restrict__ endUs, int64_t nFactors,
                                                          const double* __restrict__ noise6, double* scratch, double* records, double* cov,
                                                          double* omegaAtEnd, int* status) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= nFactors) return;
  const int64_t off = calibOff[k];
  if (off < 0) return;
  double calib[32];
#pragma unroll
  for (int i = 0; i < 32; i++) calib[i] = varData[off + i];
  const double gv[3] = {noise6[0], noise6[1], noise6[2]}, av[3] = {noise6[3], noise6[4], noise6[5]};
  const int st = pi::preintegrate(timeNs, gyro, accel, nSamples, calib, calibFlags[k], gv, av, startUs[k], endUs[k],
                                  scratch + k * (9 * pi::kMaxN), records + k * pi::kRec, cov + k * 81, omegaAtEnd + k * 3);
  if (st != pi::kOk) atomicMax(status, st);
}

void launchPreintRefresh(const int64_t* timeNs, const float* gyro, const float* accel, int64_t nSamples, const double* varData,
                         const int64_t* calibOff, const uint32_t* calibFlags, const int64_t* startUs, const int64_t* endUs, int64_t nFactors,
                         const double* noise6, double* scratch, double* records, double* cov, double* omegaAtEnd, int* status, cudaStream_t st) {
  if (nFactors <= 0) return;
  const int block = 64;
  preintRefreshKernel<<<(unsigned)((nFactors + block - 1) / block), block, 0, st>>>(timeNs, gyro, accel, nSamples, varData, calibOff, calibFlags,
                                                                                 startUs, endUs, nFactors, noise6, scratch, records, cov,
                                                                                 omegaAtEnd, status);
}

}  // namespace vb
