// Test-only: compiles the library's device routine for the rolling-shutter refresher
// (visual_inertial_bundle_adjustment_b200/csrc/dev_refresh.cuh, normally run by rsRefreshKernel) as HOST
// code so that tests/test_rs_refresh.py can compare it with the oracle without a GPU.  Not part of the
// library and never shipped: the product only runs this routine inside its CUDA kernel.
#include "../../visual_inertial_bundle_adjustment_b200/csrc/dev_refresh.cuh"

extern "C" int rf_refresh_rigs(const int64_t* timeNs, const float* gyro, const float* accel, int64_t nSamples, const double* varData,
                               const int64_t* calibOff, const int64_t* midUs, int64_t halfUs, int64_t nRigs, double* tables) {
  int worst = 0;
  for (int64_t r = 0; r < nRigs; r++) {
    if (calibOff[r] < 0) continue;
    const int st = vb::rf::refreshRig(timeNs, gyro, accel, nSamples, varData + calibOff[r], midUs[r], halfUs, tables + r * vb::rf::kStride);
    if (st > worst) worst = st;
  }
  return worst;
}
