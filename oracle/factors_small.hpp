// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// Random-walk, factory-prior and omega-prior functors:
// viba/problem/RandomWalkFactor.cpp:16-168, PriorFactor.cpp:82-176, OmegaPriorFactor.cpp:16-62.
// Per-factor constants (`c`): RW_* = diagSqrtH (23 / 17 / 6); PRIOR_IMUCALIB = diagSqrtH(23),
// prior(32); PRIOR_CAM = diagSqrtH(17), prior(17); PRIOR_SE3 = diagSqrtH(6), priorInv(7);
// OMEGA_PRIOR(_SEC) = measured omega (3).
#pragma once
#include "problem.hpp"

namespace orc {

constexpr double kOmegaPriorStd = 10.0 * M_PI / 180; // viba/common/Constants.h:19

inline int camTangentDim(uint32_t flags) {
  return 15 + ((flags & VB_CAM_EST_READOUT) ? 1 : 0) + ((flags & VB_CAM_EST_TIMEOFFSET) ? 1 : 0);
}
// VarSpec<CameraModelParam>::boxMinus — CameraModelParam.cpp:69-85
inline void camBoxMinus(const double* v, const double* base, uint32_t flags, double* d) {
  int n = 15;
  for (int i = 0; i < 15; i++) d[i] = v[i] - base[i];
  if (flags & VB_CAM_EST_READOUT) d[n++] = v[15] - base[15];
  if (flags & VB_CAM_EST_TIMEOFFSET) d[n++] = v[16] - base[16];
}

// diag-weighted boxMinus residual with E padded rows; Jp = -diag, Jn = +diag (E x tg)
inline void diagResidual(const double* delta, const double* D, int tg, int E, double* r, double* Jprev, double* Jnext) {
  for (int i = 0; i < E; i++) r[i] = 0;
  for (int i = 0; i < tg; i++) r[i] = delta[i] * D[i];
  if (Jprev) {
    for (int i = 0; i < E * tg; i++) Jprev[i] = 0;
    for (int i = 0; i < tg; i++) Jprev[i * tg + i] = -D[i];
  }
  if (Jnext) {
    for (int i = 0; i < E * tg; i++) Jnext[i] = 0;
    for (int i = 0; i < tg; i++) Jnext[i * tg + i] = D[i];
  }
}

// RandomWalkFactor.cpp:24-46
inline void rwImuCalib(const double* prev, const double* next, uint32_t flags, int tg, const double* D, double* r, double** J) {
  double d[23] = {0};
  imuCalibBoxMinus(next, prev, flags, d);
  diagResidual(d, D, tg, 23, r, J ? J[0] : nullptr, J ? J[1] : nullptr);
}
// RandomWalkFactor.cpp:69-89
inline void rwCam(const double* prev, const double* next, uint32_t flags, int tg, const double* D, double* r, double** J) {
  double d[17] = {0};
  camBoxMinus(next, prev, flags, d);
  diagResidual(d, D, tg, 17, r, J ? J[0] : nullptr, J ? J[1] : nullptr);
}
// RandomWalkFactor.cpp:107-123,141-157
inline void rwSE3(const double* prev, const double* next, const double* D, double* r, double** J) {
  double pi[7], E[7], lg[6];
  se3Inv(prev, pi);
  se3Mul(next, pi, E);
  se3Log(E, lg);
  for (int i = 0; i < 6; i++) r[i] = lg[i] * D[i];
  if (!J) return;
  double Jli[36];
  se3LeftJacobianInverse(lg, Jli);
  if (J[0]) {
    double A[36], M[36];
    se3Adj(E, A);
    matmul(Jli, A, M, 6, 6, 6);
    for (int i = 0; i < 6; i++)
      for (int j = 0; j < 6; j++) J[0][i * 6 + j] = -D[i] * M[i * 6 + j];
  }
  if (J[1])
    for (int i = 0; i < 6; i++)
      for (int j = 0; j < 6; j++) J[1][i * 6 + j] = D[i] * Jli[i * 6 + j];
}
// PriorFactor.cpp:92-107
inline void priorImuCalib(const double* x, uint32_t flags, int tg, const double* c, double* r, double** J) {
  double d[23] = {0};
  imuCalibBoxMinus(x, c + 23, flags, d);
  diagResidual(d, c, tg, 23, r, nullptr, J ? J[0] : nullptr);
}
// PriorFactor.cpp:123-136
inline void priorCam(const double* x, uint32_t flags, int tg, const double* c, double* r, double** J) {
  double d[17] = {0};
  camBoxMinus(x, c + 17, flags, d);
  diagResidual(d, c, tg, 17, r, nullptr, J ? J[0] : nullptr);
}
// PriorFactor.cpp:148-155,167-174
inline void priorSE3(const double* x, const double* c, double* r, double** J) {
  double E[7], lg[6];
  se3Mul(x, c + 6, E);
  se3Log(E, lg);
  for (int i = 0; i < 6; i++) r[i] = lg[i] * c[i];
  if (J && J[0]) {
    double Jli[36];
    se3LeftJacobianInverse(lg, Jli);
    for (int i = 0; i < 6; i++)
      for (int j = 0; j < 6; j++) J[0][i * 6 + j] = c[i] * Jli[i * 6 + j];
  }
}
// OmegaPriorFactor.cpp:24-31
inline void omegaPrior(const double* w, const double* c, double* r, double** J) {
  for (int i = 0; i < 3; i++) r[i] = (w[i] - c[i]) / kOmegaPriorStd;
  if (J && J[0]) {
    for (int i = 0; i < 9; i++) J[0][i] = 0;
    for (int i = 0; i < 3; i++) J[0][i * 3 + i] = 1.0 / kOmegaPriorStd;
  }
}
// OmegaPriorFactor.cpp:35-53
inline void omegaPriorSec(const double* w, const double* T, const double* c, double* r, double** J) {
  double R[9], wb[3];
  quatToMat(T, R);
  mat3Tvec(R, c, wb); // so3().inverse() * omega_meas
  for (int i = 0; i < 3; i++) r[i] = (w[i] - wb[i]) / kOmegaPriorStd;
  if (!J) return;
  if (J[0]) {
    for (int i = 0; i < 9; i++) J[0][i] = 0;
    for (int i = 0; i < 3; i++) J[0][i * 3 + i] = 1.0 / kOmegaPriorStd;
  }
  if (J[1]) { // [0, -R^T * hat(omega_meas) / std]
    double H[9], Rt[9], M[9];
    hat3(c, H); // -hat(-x) = hat(x)
    transpose(R, Rt, 3, 3);
    matmul(Rt, H, M, 3, 3, 3);
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) { J[1][i * 6 + j] = 0; J[1][i * 6 + 3 + j] = -M[i * 3 + j] / kOmegaPriorStd; }
  }
}

} // namespace orc
