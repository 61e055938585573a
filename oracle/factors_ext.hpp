// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// RollingShutterVisualFactor (viba/problem/VisualFactor.cpp:122-214) and
// SecondaryImuInertialFactor (viba/problem/InertialFactor.cpp:131-301).
// RS store: consts per factor = z(2), sqrtH(4), rig table index, image height; store aux =
// gravityWorld(3), number of tables, then kRsStride doubles per rig table (motion.hpp).
#pragma once
#include "factors.hpp"
#include "motion.hpp"

namespace orc {

// vars: X, T_bodyImu_world (mid exposure), T_Cam_BodyImu, cam (15 + readout + time offset), vel_world
inline bool rsVisualFactor(const double* X, const double* Tbw, const double* Tcb, const double* cam, uint32_t camFlags,
                           int camDim, const double* vel, const double* c, const double* aux, double* r, double** J) {
  const double* grav = aux;
  const double* table = aux + 4 + (int64_t)c[6] * kRsStride;
  const double tpf = c[1] / c[7] - 0.5;             // imageRow / imageHeight - 0.5
  const double dt = cam[15] * tpf - cam[16];
  const RsEstimate est = rsGetEstimate(table, grav, dt, vel, Tbw);
  double TatT_mid[7], TatT_w[7];
  se3Inv(est.T_mid_atT, TatT_mid);
  se3Mul(TatT_mid, Tbw, TatT_w);
  const bool estTime = (camFlags & (VB_CAM_EST_READOUT | VB_CAM_EST_TIMEOFFSET)) != 0;
  const bool needPoseJ = J && ((J[3] && estTime) || J[1] || J[4]);
  double JposeAtT[12];
  double* Jin[4] = {J ? J[0] : nullptr, needPoseJ ? JposeAtT : nullptr, J ? J[2] : nullptr, J ? J[3] : nullptr};
  if (!visualFactor(X, TatT_w, Tcb, cam, c, camDim, r, J ? Jin : nullptr)) return false;
  if (!J) return true;
  if (J[3] && estTime) { // numeric time derivative of the pose (VisualFactor.cpp:172-180)
    const double kEps = 1e-6;
    const RsEstimate estP = rsGetEstimate(table, grav, dt + kEps, vel, Tbw);
    double inv[7], rel[7], lg[6];
    se3Inv(estP.T_mid_atT, inv);
    se3Mul(inv, est.T_mid_atT, rel);
    se3Log(rel, lg);
    double dErr[2] = {0, 0};
    for (int e = 0; e < 2; e++) for (int k = 0; k < 6; k++) dErr[e] += JposeAtT[e * 6 + k] * (lg[k] / kEps);
    int idx = camDim;
    if (camFlags & VB_CAM_EST_TIMEOFFSET) { --idx; for (int e = 0; e < 2; e++) J[3][e * camDim + idx] = -dErr[e]; }
    if (camFlags & VB_CAM_EST_READOUT) { --idx; for (int e = 0; e < 2; e++) J[3][e * camDim + idx] = dErr[e] * tpf; }
  }
  double Rw[9];
  quatToMat(TatT_w, Rw);
  if (J[1]) {
    double A[36], v[3], Rv[3], H[9];
    se3Adj(TatT_mid, A);
    for (int i = 0; i < 3; i++) v[i] = vel[i] * dt + 0.5 * dt * dt * grav[i];
    mat3vec(Rw, v, Rv);
    hat3(Rv, H);
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) A[i * 6 + 3 + j] += H[i * 3 + j];
    matmul(JposeAtT, A, J[1], 2, 6, 6);
  }
  if (J[4]) {
    for (int e = 0; e < 2; e++)
      for (int j = 0; j < 3; j++) {
        double s = 0;
        for (int k = 0; k < 3; k++) s += JposeAtT[e * 6 + k] * Rw[k * 3 + j];
        J[4][e * 3 + j] = -dt * s;
      }
  }
  return true;
}

// SecondaryState (InertialFactor.cpp:136-188)
struct SecondaryState {
  double t_b_i[3], vel_b[3], Rwb[9], T_imu_world[7], vel_world[3];
  SecondaryState(const double* Tbw, const double* velw, const double* omega, const double* Tib) {
    double inv[7];
    se3Inv(Tib, inv);
    for (int i = 0; i < 3; i++) t_b_i[i] = inv[4 + i];
    cross3(omega, t_b_i, vel_b);
    double Rbw[9];
    quatToMat(Tbw, Rbw);
    transpose(Rbw, Rwb, 3, 3);
    se3Mul(Tib, Tbw, T_imu_world);
    double t[3];
    mat3vec(Rwb, vel_b, t);
    for (int i = 0; i < 3; i++) vel_world[i] = velw[i] + t[i];
  }
  // Jp 9x6 (T_imu_world), Jv 9x3 (imu_vel_world) -> outputs (any may be null)
  void compose(const double* omega, const double* Tib, const double* Jp, const double* Jv, double* JT, double* JV, double* JW, double* JE) const {
    if (JT) {
      double A[36], H[9], RH[9], M[54];
      se3Adj(Tib, A);
      matmul(Jp, A, M, 9, 6, 6);
      hat3(vel_b, H);                         // -hat(-v) = hat(v)
      matmul(Rwb, H, RH, 3, 3, 3);
      for (int i = 0; i < 9; i++)
        for (int j = 0; j < 3; j++) {
          double s = 0;
          for (int k = 0; k < 3; k++) s += Jv[i * 3 + k] * RH[k * 3 + j];
          M[i * 6 + 3 + j] += s;
        }
      for (int i = 0; i < 54; i++) JT[i] = M[i];
    }
    if (JV) for (int i = 0; i < 27; i++) JV[i] = Jv[i];
    if (JW) {
      double nt[3] = {-t_b_i[0], -t_b_i[1], -t_b_i[2]}, H[9], RH[9];
      hat3(nt, H);
      matmul(Rwb, H, RH, 3, 3, 3);
      matmul(Jv, RH, JW, 9, 3, 3);
    }
    if (JE) {
      double H[9], Rib[9], RibT[9], RH[9], M[9];
      hat3(omega, H);
      quatToMat(Tib, Rib);                    // so3().Adj() = R ; (-R^T)
      matmul(Rwb, H, RH, 3, 3, 3);
      transpose(Rib, RibT, 3, 3);
      matmul(RH, RibT, M, 3, 3, 3);           // R_wb hat(w) R_ib^T ; the block is -M

      for (int i = 0; i < 9; i++)
        for (int j = 0; j < 6; j++) {
          double s = Jp[i * 6 + j];
          if (j < 3) for (int k = 0; k < 3; k++) s -= Jv[i * 3 + k] * M[k * 3 + j];
          JE[i * 6 + j] = s;
        }
    }
  }
};

// split extrinsics: vars calib, Tp, vp, wp, Ep, Tn, vn, wn, En, g  (InertialFactor.cpp:190-254)
inline void secondaryInertialFactor(const double* calib, uint32_t cflags, int cdim, const double* Tp, const double* vp,
                                    const double* wp, const double* Ep, const double* Tn, const double* vn, const double* wn,
                                    const double* En, const double* g, const double* pre, double* r, double** J) {
  const SecondaryState ps(Tp, vp, wp, Ep), ns(Tn, vn, wn, En);
  double JpP[54], JvP[27], JpN[54], JvN[27];
  double* Jin[6] = {J ? J[0] : nullptr, JpP, JvP, JpN, JvN, J ? J[9] : nullptr};
  inertialFactor(calib, cflags, cdim, ps.T_imu_world, ps.vel_world, ns.T_imu_world, ns.vel_world, g, pre, r, J ? Jin : nullptr);
  if (!J) return;
  ps.compose(wp, Ep, JpP, JvP, J[1], J[2], J[3], J[4]);
  ns.compose(wn, En, JpN, JvN, J[5], J[6], J[7], J[8]);
}
// common extrinsics: vars calib, Tp, vp, wp, Tn, vn, wn, E, g  (InertialFactor.cpp:256-301)
inline void secondaryInertialFactorCommon(const double* calib, uint32_t cflags, int cdim, const double* Tp, const double* vp,
                                          const double* wp, const double* Tn, const double* vn, const double* wn,
                                          const double* E, const double* g, const double* pre, double* r, double** J) {
  double JeP[54], JeN[54];
  double* Jin[10] = {nullptr};
  if (J) { Jin[0] = J[0]; Jin[1] = J[1]; Jin[2] = J[2]; Jin[3] = J[3]; Jin[4] = J[7] ? JeP : nullptr;
           Jin[5] = J[4]; Jin[6] = J[5]; Jin[7] = J[6]; Jin[8] = J[7] ? JeN : nullptr; Jin[9] = J[8]; }
  secondaryInertialFactor(calib, cflags, cdim, Tp, vp, wp, E, Tn, vn, wn, E, g, pre, r, J ? Jin : nullptr);
  if (J && J[7]) for (int i = 0; i < 54; i++) J[7][i] = JeP[i] + JeN[i];
}

} // namespace orc
