// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// Rot/vel/pos motion integrals used by the rolling-shutter visual factor:
// lib/motion/preintegration/MotionIntegral.cpp:28-160 (combine, uncombineLeft, differentiate,
// integrate; Q6: F6 = 729.0 is kept as in the reference) and
// lib/motion/preintegration/RollingShutterData.cpp:67-111 (getEstimate).
// Per-rig table layout (kRsStride doubles): [n, samples 32 x (q4 dV3 dP3 dt), interp 31 x (gyro3 accel3 dVel3)].
#pragma once
#include <stdexcept>
#include <string>
#include "vmath.hpp"

namespace orc {

constexpr int kRsMaxSamples = 32;
constexpr int kRsStride = 1 + kRsMaxSamples * 11 + (kRsMaxSamples - 1) * 9;
constexpr int kRsInterpOff = 1 + kRsMaxSamples * 11;

struct Rvp { double q[4], dV[3], dP[3], dt; };

inline void quatRot(const double* q, const double* v, double* o) { double R[9]; quatToMat(q, R); mat3vec(R, v, o); }

// MotionIntegral.cpp:28-33
inline Rvp combine(const Rvp& a, const Rvp& b) {
  Rvp c;
  quatMul(a.q, b.q, c.q);
  double t[3];
  quatRot(a.q, b.dV, t);
  for (int i = 0; i < 3; i++) c.dV[i] = a.dV[i] + t[i];
  quatRot(a.q, b.dP, t);
  for (int i = 0; i < 3; i++) c.dP[i] = a.dP[i] + a.dV[i] * b.dt + t[i];
  c.dt = a.dt + b.dt;
  return c;
}
// MotionIntegral.cpp:117-160
inline Rvp integrate(const double* gyro, const double* accel, double dt) {
  double om[3], up[3];
  for (int i = 0; i < 3; i++) { om[i] = gyro[i] * dt; up[i] = accel[i] * dt; }
  Rvp r;
  so3Exp(om, r.q);
  const double th2 = dot3(om, om), th = std::sqrt(th2), th4 = th2 * th2;
  double c1, c2, c3;
  if (th < 1e-3) {
    c1 = (1.0 / 2.0) - (th2 / 24.0) + (th4 / 729.0); // F6 = 729 (sic), MotionIntegral.cpp:82
    c2 = (1.0 / 6.0) - (th2 / 120.0) + (th4 / 5040.0);
    c3 = (1.0 / 24.0) - (th2 / 729.0) + (th4 / 40320.0);
  } else {
    const double s = std::sin(th) / th, m = (1.0 - std::cos(th)) / th2;
    c1 = m; c2 = (1.0 - s) / th2; c3 = (0.5 - m) / th2;
  }
  double W[9], W2[9];
  hat3(om, W);
  matmul(W, W, W2, 3, 3, 3);
  double U2V[9], U2P[9];
  for (int i = 0; i < 9; i++) { U2V[i] = c1 * W[i] + c2 * W2[i]; U2P[i] = c2 * W[i] + c3 * W2[i]; }
  U2V[0] += 1; U2V[4] += 1; U2V[8] += 1;
  U2P[0] += 0.5; U2P[4] += 0.5; U2P[8] += 0.5;
  mat3vec(U2V, up, r.dV);
  double upd[3] = {up[0] * dt, up[1] * dt, up[2] * dt};
  mat3vec(U2P, upd, r.dP);
  r.dt = dt;
  return r;
}
// MotionIntegral.cpp:111-115: interp = gyro(3) accel(3) deltaVel(3)
inline Rvp integrateInterp(const double* interp, double dt) {
  Rvp r = integrate(interp, interp + 3, dt);
  for (int i = 0; i < 3; i++) r.dP[i] += interp[6 + i] * dt;
  return r;
}

struct RsEstimate { double T_mid_atT[7]; };

// RollingShutterData::getEstimate (RollingShutterData.cpp:67-111); T_bw = T_bodyImu_world at mid
// exposure (the factor passes its inverse and getEstimate inverts it back)
inline RsEstimate rsGetEstimate(const double* table, const double* gravityWorld, double tDelta, const double* velWorld, const double* Tbw) {
  const int n = (int)table[0];
  if (n <= 0) throw std::runtime_error("Not initialized RS?");
  const double* smp = table + 1;
  int idx = 0; // first sample with dt > tDelta
  while (idx < n && !(tDelta < smp[idx * 11 + 10])) idx++;
  if (idx == n || idx == 0) throw std::runtime_error("RollingShutterData::getEstimate: out of range");
  Rvp prev;
  const double* s = smp + (idx - 1) * 11;
  for (int i = 0; i < 4; i++) prev.q[i] = s[i];
  for (int i = 0; i < 3; i++) { prev.dV[i] = s[4 + i]; prev.dP[i] = s[7 + i]; }
  prev.dt = s[10];
  const Rvp at = combine(prev, integrateInterp(table + kRsInterpOff + (idx - 1) * 9, tDelta - prev.dt));
  double gm[3], vm[3];
  quatRot(Tbw, gravityWorld, gm);
  quatRot(Tbw, velWorld, vm);
  RsEstimate e;
  for (int i = 0; i < 4; i++) e.T_mid_atT[i] = at.q[i];
  for (int i = 0; i < 3; i++) e.T_mid_atT[4 + i] = at.dP[i] + vm[i] * tDelta + gm[i] * (0.5 * tDelta * tDelta);
  return e;
}

} // namespace orc
