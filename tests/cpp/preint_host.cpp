// Test-only: compiles the library's device routine for the preintegration refresher
// (visual_inertial_bundle_adjustment_b200/csrc/dev_preint.cuh, normally run by preintRefreshKernel) as HOST code so
// that tests/test_preint_device.py can compare it with the oracle without a GPU.  Not part of the library and never
// shipped: the product only runs this routine inside its CUDA kernel.
#include "../../visual_inertial_bundle_adjustment_b200/csrc/dev_preint.cuh"
#include <vector>

extern "C" int pi_preintegrate(const int64_t* timeNs, const float* gyro, const float* accel, int64_t nSamples, const double* calib32,
                               uint32_t flags, const double* gyroVar, const double* accelVar, int64_t startUs, int64_t endUs, double* rec250,
                               double* cov81, double* omega3) {
  std::vector<double> J(9 * vb::pi::kMaxN);
  return vb::pi::preintegrate(timeNs, gyro, accel, nSamples, calib32, flags, gyroVar, accelVar, startUs, endUs, J.data(), rec250, cov81, omega3);
}
