// Rolling-shutter table refresher: one thread per rig runs RollingShutterData::compute
// (dev_refresh.cuh) on the raw IMU-0 samples with the current calibration of the rig's window and writes
// the rig's table in place.  ~30 sequential integration steps per rig, a few thousand rigs: the kernel is
// latency-sized (tens of microseconds), its job is to keep the host and its 2 x 8 B x 632 x rigs of
// table traffic out of the iteration.
#include "dev_refresh.cuh"
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

}  // namespace vb
