// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// Residual + Jacobian functors, restating viba/problem/*.cpp one by one.
// Jacobians are row-major E x d_i; a null J[i] means "not requested" (constant variable or
// cost-only evaluation), like the NullRef convention of lib/small_thing/Common.h:123-204.
#pragma once
#include "problem.hpp"
#include "camera.hpp"

namespace orc {

constexpr int kMaxE = 23;     // padded IMU-calibration RW/prior residual (RandomWalkFactor.cpp:29)
constexpr int kMaxVars = 10;  // SecondaryImuInertialFactor split extrinsics (InertialFactor.cpp:190-254)
constexpr int kPreintStride = 250; // R(4) dV(3) dP(3) dt(1) J(9x23) evalPoint(32)

inline int factorErrorSize(int type) {
  switch (type) {
    case VB_F_VISUAL: case VB_F_VISUAL_RS: return 2;
    case VB_F_INERTIAL: case VB_F_INERTIAL_SEC: case VB_F_INERTIAL_SEC_SPLIT: return 9;
    case VB_F_RW_IMUCALIB: case VB_F_PRIOR_IMUCALIB: return 23;
    case VB_F_RW_CAM: case VB_F_PRIOR_CAM: return 17;
    case VB_F_RW_SE3: case VB_F_PRIOR_SE3: return 6;
    case VB_F_OMEGA_PRIOR: case VB_F_OMEGA_PRIOR_SEC: return 3;
  }
  return 0;
}
inline bool factorHasPrecision(int type) {
  return type == VB_F_INERTIAL || type == VB_F_INERTIAL_SEC || type == VB_F_INERTIAL_SEC_SPLIT;
}

// VisualFactor::operator() — viba/problem/VisualFactor.cpp:40-82
inline bool visualFactor(const double* X, const double* Tbw, const double* Tcb, const double* cam,
                         const double* c /* z(2), sqrtH(4) */, int camDim, double* r, double** J) {
  double pr[3], pc[3];
  se3Act(Tbw, X, pr);
  se3Act(Tcb, pr, pc);
  double uv[2], dP[6], dC[30];
  const bool needJ = J && (J[0] || J[1] || J[2] || J[3]);
  if (!fisheye624Project(pc, cam, uv, needJ ? dP : nullptr, (J && J[3]) ? dC : nullptr)) return false;
  const double e0 = uv[0] - c[0], e1 = uv[1] - c[1];
  const double* S = c + 2;
  r[0] = S[0] * e0 + S[1] * e1;
  r[1] = S[2] * e0 + S[3] * e1;
  if (!needJ) return true;
  double W[6]; // sqrtH * dProj_dPointCam  (2x3)
  matmul(S, dP, W, 2, 2, 3);
  double Rcb[9], Rbw[9];
  quatToMat(Tcb, Rcb);
  quatToMat(Tbw, Rbw);
  if (J[0]) {
    double RR[9];
    matmul(Rcb, Rbw, RR, 3, 3, 3);
    matmul(W, RR, J[0], 2, 3, 3);
  }
  if (J[1]) {
    double A[6], H[9], AH[6];
    matmul(W, Rcb, A, 2, 3, 3);
    double npr[3] = {-pr[0], -pr[1], -pr[2]};
    hat3(npr, H);
    matmul(A, H, AH, 2, 3, 3);
    for (int i = 0; i < 2; i++)
      for (int j = 0; j < 3; j++) { J[1][i * 6 + j] = A[i * 3 + j]; J[1][i * 6 + 3 + j] = AH[i * 3 + j]; }
  }
  if (J[2]) {
    double H[9], WH[6];
    double npc[3] = {-pc[0], -pc[1], -pc[2]};
    hat3(npc, H);
    matmul(W, H, WH, 2, 3, 3);
    for (int i = 0; i < 2; i++)
      for (int j = 0; j < 3; j++) { J[2][i * 6 + j] = W[i * 3 + j]; J[2][i * 6 + 3 + j] = WH[i * 3 + j]; }
  }
  if (J[3]) { // leftCols(15) = sqrtH * dProj_dCamParams ; remaining columns (none for B1) zero
    for (int i = 0; i < 2 * camDim; i++) J[3][i] = 0;
    for (int j = 0; j < 15; j++) {
      J[3][j] = S[0] * dC[j] + S[1] * dC[15 + j];
      J[3][camDim + j] = S[2] * dC[j] + S[3] * dC[15 + j];
    }
  }
  return true;
}

// InertialFactor::operator() — viba/problem/InertialFactor.cpp:23-123
// vars: calib, Tp, vp, Tn, vn, gravity ; pre = preintegration record (kPreintStride doubles)
inline void inertialFactor(const double* calib, uint32_t cflags, int cdim, const double* Tp, const double* vp,
                           const double* Tn, const double* vn, const double* g, const double* pre,
                           double* r, double** J) {
  const double* preR = pre;
  const double* dV = pre + 4;
  const double* dP = pre + 7;
  const double dt = pre[10];
  const double* PJ = pre + 11;       // 9 x 23 row-major, first cdim columns used
  const double* evalPoint = pre + 218;

  double dc[23] = {0};
  imuCalibBoxMinus(calib, evalPoint, cflags, dc);
  double corr[9];
  for (int i = 0; i < 9; i++) {
    double s = 0;
    for (int k = 0; k < cdim; k++) s += PJ[i * 23 + k] * dc[k];
    corr[i] = s;
  }
  // rotation error
  double nc[3] = {-corr[0], -corr[1], -corr[2]};
  double qc[4], qRi[4], qCorr[4], qn_inv[4], t1[4], qErr[4];
  so3Exp(nc, qc);
  quatConj(preR, qRi);
  quatMul(qc, qRi, qCorr);          // corrected_R_ImuNext_ImuPrev
  quatConj(Tn, qn_inv);
  quatMul(qCorr, Tp, t1);
  quatMul(t1, qn_inv, qErr);        // R_rotErr
  double lg[3];
  so3Log(qErr, lg);
  double logRotErr[3] = {-lg[0], -lg[1], -lg[2]};

  double Rp[9];
  quatToMat(Tp, Rp);
  double dvw[3], dvp[3];
  for (int i = 0; i < 3; i++) dvw[i] = vn[i] - vp[i] - g[i] * dt;
  mat3vec(Rp, dvw, dvp);            // deltaVelocity_prevImu_est
  double velErr[3];
  for (int i = 0; i < 3; i++) velErr[i] = dV[i] - dvp[i] + corr[3 + i];

  double qpn[4], Rpn[9];
  quatMul(Tp, qn_inv, qpn);         // R_prevImu_nextImu
  quatToMat(qpn, Rpn);
  double a[3], b[3], tmp[3], dpp[3];
  mat3vec(Rpn, Tn + 4, a);
  for (int i = 0; i < 3; i++) tmp[i] = vp[i] * dt + g[i] * (0.5 * dt * dt);
  mat3vec(Rp, tmp, b);
  for (int i = 0; i < 3; i++) dpp[i] = Tp[4 + i] - a[i] - b[i]; // deltaPosition_prevImu_est
  double posErr[3];
  for (int i = 0; i < 3; i++) posErr[i] = dP[i] - dpp[i] + corr[6 + i];

  for (int i = 0; i < 3; i++) { r[i] = logRotErr[i]; r[3 + i] = velErr[i]; r[6 + i] = posErr[i]; }
  if (!J) return;

  double nl[3] = {-logRotErr[0], -logRotErr[1], -logRotErr[2]};
  double Jli[9];
  so3LeftJacobianInverse(nl, Jli);  // dLogRotErr_dLeftRotError
  if (J[1]) { // prev pose 9x6
    double Rc[9], M[9], Hv[9], Hp[9];
    quatToMat(qCorr, Rc);
    matmul(Jli, Rc, M, 3, 3, 3);
    hat3(dvp, Hv);                  // -hat(-x) = hat(x)
    hat3(dpp, Hp);
    for (int i = 0; i < 54; i++) J[1][i] = 0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) {
        J[1][i * 6 + 3 + j] = -M[i * 3 + j];
        J[1][(3 + i) * 6 + 3 + j] = Hv[i * 3 + j];
        J[1][(6 + i) * 6 + 3 + j] = Hp[i * 3 + j];
      }
    for (int i = 0; i < 3; i++) J[1][(6 + i) * 6 + i] = -1.0;
  }
  if (J[2]) { // prev vel 9x3
    for (int i = 0; i < 27; i++) J[2][i] = 0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) {
        J[2][(3 + i) * 3 + j] = Rp[i * 3 + j];
        J[2][(6 + i) * 3 + j] = Rp[i * 3 + j] * dt;
      }
  }
  if (J[3]) { // next pose 9x6
    double Re[9], M[9];
    quatToMat(qErr, Re);
    matmul(Jli, Re, M, 3, 3, 3);
    for (int i = 0; i < 54; i++) J[3][i] = 0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) {
        J[3][i * 6 + 3 + j] = M[i * 3 + j];
        J[3][(6 + i) * 6 + j] = Rpn[i * 3 + j];
      }
  }
  if (J[4]) { // next vel 9x3
    for (int i = 0; i < 27; i++) J[4][i] = 0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) J[4][(3 + i) * 3 + j] = -Rp[i * 3 + j];
  }
  if (J[5]) { // gravity 9x2
    double O[6], Ot[6], RO[6];
    s2Ortho(g, O);
    transpose(O, Ot, 2, 3);
    matmul(Rp, Ot, RO, 3, 3, 2);
    for (int i = 0; i < 18; i++) J[5][i] = 0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 2; j++) {
        J[5][(3 + i) * 2 + j] = dt * RO[i * 2 + j];
        J[5][(6 + i) * 2 + j] = dt * RO[i * 2 + j] * (0.5 * dt);
      }
  }
  if (J[0]) { // calib 9 x cdim
    double Jl[9], M[9];
    so3LeftJacobian(nc, Jl);
    matmul(Jli, Jl, M, 3, 3, 3);    // dLogRotErr_dResidualCorrection
    for (int i = 0; i < 3; i++)
      for (int k = 0; k < cdim; k++) {
        double s = 0;
        for (int l = 0; l < 3; l++) s += M[i * 3 + l] * PJ[l * 23 + k];
        J[0][i * cdim + k] = s;
      }
    for (int i = 3; i < 9; i++)
      for (int k = 0; k < cdim; k++) J[0][i * cdim + k] = PJ[i * 23 + k];
  }
}

} // namespace orc
