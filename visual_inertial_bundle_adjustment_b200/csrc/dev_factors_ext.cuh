// Device routines for the rolling-shutter visual factor (viba/problem/VisualFactor.cpp:122-214 with
// RollingShutterData::getEstimate, lib/motion/preintegration/RollingShutterData.cpp:67-111, and the
// motion integrals of MotionIntegral.cpp:28-33,111-160) and for the secondary-IMU inertial factor
// (viba/problem/InertialFactor.cpp:131-301).
// RS store: consts per factor = z(2), sqrtH(4), rig table index, image height; store aux =
// gravityWorld(3), table count, then kRsStride doubles per rig:
// [n, samples 32 x (q4 dV3 dP3 dt), interp 31 x (gyro3 accel3 dVel3)].
#pragma once
#include "dev_factors_small.cuh"

namespace vb {

constexpr int kRsMaxSamples = 32;
constexpr int kRsStride = 1 + kRsMaxSamples * 11 + (kRsMaxSamples - 1) * 9;
constexpr int kRsInterpOff = 1 + kRsMaxSamples * 11;

VB_DEV void qrot(const double* q, const double* v, double* o) { double R[9]; q2R(q, R); m3v(R, v, o); }

// integrate(interp, dt): rotation quaternion q, dV, dP (MotionIntegral.cpp:111-160; F6 = 729 as there)
VB_DEV void rvp_integrate(const double* interp, double dt, double* q, double* dV, double* dP) {
  double om[3], up[3];
#pragma unroll
  for (int i = 0; i < 3; i++) { om[i] = interp[i] * dt; up[i] = interp[3 + i] * dt; }
  so3_exp(om, q);
  const double th2 = dot3(om, om), th = sqrt(th2), th4 = th2 * th2;
  double c1, c2, c3;
  if (th < 1e-3) {
    c1 = 0.5 - th2 / 24.0 + th4 / 729.0;
    c2 = 1.0 / 6.0 - th2 / 120.0 + th4 / 5040.0;
    c3 = 1.0 / 24.0 - th2 / 729.0 + th4 / 40320.0;
  } else {
    double s, c;
    sincos(th, &s, &c);
    const double m = (1.0 - c) / th2;
    c1 = m; c2 = (1.0 - s / th) / th2; c3 = (0.5 - m) / th2;
  }
  double a[3], b[3], upd[3];
  cross3(om, up, a); cross3(om, a, b);
#pragma unroll
  for (int i = 0; i < 3; i++) { dV[i] = up[i] + c1 * a[i] + c2 * b[i]; upd[i] = up[i] * dt; }
  cross3(om, upd, a); cross3(om, a, b);
#pragma unroll
  for (int i = 0; i < 3; i++) dP[i] = 0.5 * upd[i] + c2 * a[i] + c3 * b[i] + interp[6 + i] * dt;
}

// T_midImu_imuAtT (7) from the rig table; returns false when tDelta leaves the sampled window
// (the reference throws there: RollingShutterData.cpp:83-91)
VB_DEV bool rs_estimate(const double* table, const double* grav, double tDelta, const double* velW, const double* Tbw, double* Tmid_atT) {
  const int n = (int)table[0];
  const double* smp = table + 1;
  if (n <= 0) return false;
  // first sample with dt > tDelta (upper_bound, as the reference searches): the samples sit on the gyro boundaries,
  // almost uniformly spaced, so start from the proportional guess and walk — two independent loads and usually one
  // or two checks instead of the five dependent loads of a bisection
  const double tFirst = smp[10], tLast = smp[(n - 1) * 11 + 10];
  int lo = tLast > tFirst ? (int)((tDelta - tFirst) / (tLast - tFirst) * (double)(n - 1)) + 1 : 0;
  lo = lo < 0 ? 0 : (lo > n ? n : lo);
  while (lo > 0 && tDelta < smp[(lo - 1) * 11 + 10]) lo--;
  while (lo < n && !(tDelta < smp[lo * 11 + 10])) lo++;
  if (lo == n || lo == 0) return false;
  const double* s = smp + (lo - 1) * 11;
  double q[4], dV[3], dP[3];
  const double dtb = tDelta - s[10];
  rvp_integrate(table + kRsInterpOff + (lo - 1) * 9, dtb, q, dV, dP);
  // combine(prev, b): R = Ra Rb ; dP = a.dP + a.dV * b.dt + Ra b.dP
  qmul(s, q, Tmid_atT);
  double t[3], gm[3], vm[3];
  qrot(s, dP, t);
  qrot(Tbw, grav, gm);
  qrot(Tbw, velW, vm);
#pragma unroll
  for (int i = 0; i < 3; i++)
    Tmid_atT[4 + i] = s[7 + i] + s[4 + i] * dtb + t[i] + vm[i] * tDelta + gm[i] * (0.5 * tDelta * tDelta);
  return true;
}

// Outputs: r(2), JX 2x3, JT 2x6 (mid-exposure rig pose), JE 2x6, JC 2x17 (15 + readout + time offset
// columns; the caller maps them to the variable's tangent layout), JV 2x3.  status: 0 valid,
// 1 invalid (optional residual empty), 2 time shift out of the table (error)
template <bool WANT_J>
VB_DEV int rs_visual_factor(const double* X, const double* Tbw, const double* Tcb, const double* cam, uint32_t camFlags,
                            const double* vel, const double* c, const double* aux, double* r, double* JX, double* JT,
                            double* JE, double* JC, double* JV) {
  const double* grav = aux;
  const double* table = aux + 4 + (int64_t)c[6] * kRsStride;
  const double tpf = c[1] / c[7] - 0.5;
  const double dt = cam[15] * tpf - cam[16];
  double Tm[7];
  if (!rs_estimate(table, grav, dt, vel, Tbw, Tm)) return 2;
  double TatT_mid[7], TatT_w[7];
  se3_inv(Tm, TatT_mid);
  se3_mul(TatT_mid, Tbw, TatT_w);
  double JP[12], JC15[30];
  if (!visual_factor<WANT_J>(X, TatT_w, Tcb, cam, c, r, JX, JP, JE, JC15)) return 1;
  if (!WANT_J) return 0;
#pragma unroll
  for (int j = 0; j < 15; j++) { JC[j] = JC15[j]; JC[17 + j] = JC15[15 + j]; }
  JC[15] = JC[16] = JC[17 + 15] = JC[17 + 16] = 0.0;
  if (camFlags & (VB_CAM_EST_READOUT | VB_CAM_EST_TIMEOFFSET)) { // numeric time derivative (VisualFactor.cpp:172-180)
    const double kEps = 1e-6;
    double Tp[7], inv[7], rel[7], lg[6];
    if (!rs_estimate(table, grav, dt + kEps, vel, Tbw, Tp)) return 2;
    se3_inv(Tp, inv);
    se3_mul(inv, Tm, rel);
    se3_log(rel, lg);
    double d0 = 0, d1 = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) { d0 = fma(JP[k], lg[k] / kEps, d0); d1 = fma(JP[6 + k], lg[k] / kEps, d1); }
    if (camFlags & VB_CAM_EST_READOUT) { JC[15] = d0 * tpf; JC[17 + 15] = d1 * tpf; }
    if (camFlags & VB_CAM_EST_TIMEOFFSET) { JC[16] = -d0; JC[17 + 16] = -d1; }
  }
  double Rw[9];
  q2R(TatT_w, Rw);
  {
    double A[36], v[3], Rv[3];
    se3_adj(TatT_mid, A);
#pragma unroll
    for (int i = 0; i < 3; i++) v[i] = vel[i] * dt + 0.5 * dt * dt * grav[i];
    m3v(Rw, v, Rv);
    A[0 * 6 + 4] += -Rv[2]; A[0 * 6 + 5] += Rv[1];
    A[1 * 6 + 3] += Rv[2];  A[1 * 6 + 5] += -Rv[0];
    A[2 * 6 + 3] += -Rv[1]; A[2 * 6 + 4] += Rv[0];
    mm<2, 6, 6>(JP, A, JT);
  }
#pragma unroll
  for (int e = 0; e < 2; e++)
#pragma unroll
    for (int j = 0; j < 3; j++)
      JV[e * 3 + j] = -dt * (JP[e * 6] * Rw[j] + JP[e * 6 + 1] * Rw[3 + j] + JP[e * 6 + 2] * Rw[6 + j]);
  return 0;
}

// ---- SecondaryImuInertialFactor ---------------------------------------------------------------------
struct SecState {
  double tbi[3], velb[3], Rwb[9], Tiw[7], velw[3];
  VB_DEV SecState(const double* Tbw, const double* vw, const double* om, const double* Tib) {
    double inv[7], Rbw[9], t[3];
    se3_inv(Tib, inv);
#pragma unroll
    for (int i = 0; i < 3; i++) tbi[i] = inv[4 + i];
    cross3(om, tbi, velb);
    q2R(Tbw, Rbw);
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < 3; j++) Rwb[i * 3 + j] = Rbw[j * 3 + i];
    se3_mul(Tib, Tbw, Tiw);
    m3v(Rwb, velb, t);
#pragma unroll
    for (int i = 0; i < 3; i++) velw[i] = vw[i] + t[i];
  }
  // composeJacobians (InertialFactor.cpp:148-180): Jp 9x6, Jv 9x3 -> JT 9x6, JV 9x3, JW 9x3, JE 9x6
  VB_DEV void compose(const double* om, const double* Tib, const double* Jp, const double* Jv, double* JT, double* JV,
                      double* JW, double* JE) const {
    if (JT) {
      double A[36], H[9], RH[9];
      se3_adj(Tib, A);
      hat3(velb, H);
      mm<3, 3, 3>(Rwb, H, RH);
      for (int i = 0; i < 9; i++)
        for (int j = 0; j < 6; j++) {
          double s = 0;
          for (int k = 0; k < 6; k++) s = fma(Jp[i * 6 + k], A[k * 6 + j], s);
          if (j >= 3) for (int k = 0; k < 3; k++) s = fma(Jv[i * 3 + k], RH[k * 3 + j - 3], s);
          JT[i * 6 + j] = s;
        }
    }
    if (JV) for (int i = 0; i < 27; i++) JV[i] = Jv[i];
    if (JW) {
      const double nt[3] = {-tbi[0], -tbi[1], -tbi[2]};
      double H[9], RH[9];
      hat3(nt, H);
      mm<3, 3, 3>(Rwb, H, RH);
      for (int i = 0; i < 9; i++)
        for (int j = 0; j < 3; j++) JW[i * 3 + j] = Jv[i * 3] * RH[j] + Jv[i * 3 + 1] * RH[3 + j] + Jv[i * 3 + 2] * RH[6 + j];
    }
    if (JE) {
      double H[9], Rib[9], RH[9], M[9];
      hat3(om, H);
      q2R(Tib, Rib);
      mm<3, 3, 3>(Rwb, H, RH);
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) M[i * 3 + j] = RH[i * 3] * Rib[j * 3] + RH[i * 3 + 1] * Rib[j * 3 + 1] + RH[i * 3 + 2] * Rib[j * 3 + 2];
      for (int i = 0; i < 9; i++)
        for (int j = 0; j < 6; j++) {
          double s = Jp[i * 6 + j];
          if (j < 3) for (int k = 0; k < 3; k++) s = fma(-Jv[i * 3 + k], M[k * 3 + j], s);
          JE[i * 6 + j] = s;
        }
    }
  }
};

// split extrinsics: J order = calib, Tp, vp, wp, Ep, Tn, vn, wn, En, g
VB_DEV void secondary_inertial(const double* calib, uint32_t cflags, int cdim, const double* Tp, const double* vp,
                               const double* wp, const double* Ep, const double* Tn, const double* vn, const double* wn,
                               const double* En, const double* g, const double* pre, double* r, double* const* J) {
  const SecState ps(Tp, vp, wp, Ep), ns(Tn, vn, wn, En);
  double JpP[54], JvP[27], JpN[54], JvN[27];
  double* Jin[6] = {J ? J[0] : nullptr, JpP, JvP, JpN, JvN, J ? J[9] : nullptr};
  inertial_factor(calib, cflags, cdim, ps.Tiw, ps.velw, ns.Tiw, ns.velw, g, pre, r, J ? Jin : nullptr);
  if (!J) return;
  ps.compose(wp, Ep, JpP, JvP, J[1], J[2], J[3], J[4]);
  ns.compose(wn, En, JpN, JvN, J[5], J[6], J[7], J[8]);
}
// common extrinsics: J order = calib, Tp, vp, wp, Tn, vn, wn, E, g ; scratch: 108 doubles for the two
// extrinsics Jacobians that are summed (InertialFactor.cpp:297-299)
VB_DEV void secondary_inertial_common(const double* calib, uint32_t cflags, int cdim, const double* Tp, const double* vp,
                                      const double* wp, const double* Tn, const double* vn, const double* wn, const double* E,
                                      const double* g, const double* pre, double* r, double* const* J) {
  double JeP[54], JeN[54];
  double* Jin[10];
  if (J) { Jin[0] = J[0]; Jin[1] = J[1]; Jin[2] = J[2]; Jin[3] = J[3]; Jin[4] = J[7] ? JeP : nullptr;
           Jin[5] = J[4]; Jin[6] = J[5]; Jin[7] = J[6]; Jin[8] = J[7] ? JeN : nullptr; Jin[9] = J[8]; }
  secondary_inertial(calib, cflags, cdim, Tp, vp, wp, E, Tn, vn, wn, E, g, pre, r, J ? Jin : nullptr);
  if (J && J[7]) for (int i = 0; i < 54; i++) J[7][i] = JeP[i] + JeN[i];
}

} // namespace vb
