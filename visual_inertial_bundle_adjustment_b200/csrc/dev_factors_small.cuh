// Device residual / Jacobian routines for the low-count factor types (inertial, random walk,
// factory prior, omega prior).  One thread evaluates one factor; Jacobians go to thread-local
// buffers (row-major E x d_i), null pointer = not requested.
#pragma once
#include "dev_factors.cuh"

namespace vb {

// ---- InertialFactor::operator() — viba/problem/InertialFactor.cpp:23-123 -------------------------
// vars: calib, Tp, vp, Tn, vn, gravity ; pre = preintegration record (kPreintStride doubles):
// R_prev_next(4) dV(3) dP(3) dt(1) J(9x23 row-major) evalPoint(32)
VB_DEV void inertial_factor(const double* calib, uint32_t cflags, int cdim, const double* Tp, const double* vp,
                            const double* Tn, const double* vn, const double* g, const double* pre,
                            double* r, double* const* J) {
  const double* preR = pre;
  const double* dV = pre + 4;
  const double* dP = pre + 7;
  const double dt = pre[10];
  const double* PJ = pre + 11;
  const double* evalPoint = pre + 218;
  double dc[23];
  for (int i = 0; i < 23; i++) dc[i] = 0;
  imu_boxminus(calib, evalPoint, cflags, dc);
  double corr[9];
  for (int i = 0; i < 9; i++) {
    double s = 0;
    for (int k = 0; k < cdim; k++) s = fma(PJ[i * 23 + k], dc[k], s);
    corr[i] = s;
  }
  const double nc[3] = {-corr[0], -corr[1], -corr[2]};
  double qc[4], qRi[4], qCorr[4], qni[4], t1[4], qErr[4];
  so3_exp(nc, qc);
  qconj(preR, qRi);
  qmul(qc, qRi, qCorr);
  qconj(Tn, qni);
  qmul(qCorr, Tp, t1);
  qmul(t1, qni, qErr);
  double lg[3];
  so3_log(qErr, lg);
  double Rp[9];
  q2R(Tp, Rp);
  double dvw[3], dvp[3];
  for (int i = 0; i < 3; i++) dvw[i] = vn[i] - vp[i] - g[i] * dt;
  m3v(Rp, dvw, dvp);
  double qpn[4], Rpn[9], a[3], b[3], tmp[3], dpp[3];
  qmul(Tp, qni, qpn);
  q2R(qpn, Rpn);
  m3v(Rpn, Tn + 4, a);
  for (int i = 0; i < 3; i++) tmp[i] = vp[i] * dt + g[i] * (0.5 * dt * dt);
  m3v(Rp, tmp, b);
  for (int i = 0; i < 3; i++) dpp[i] = Tp[4 + i] - a[i] - b[i];
  for (int i = 0; i < 3; i++) {
    r[i] = -lg[i];
    r[3 + i] = dV[i] - dvp[i] + corr[3 + i];
    r[6 + i] = dP[i] - dpp[i] + corr[6 + i];
  }
  if (!J) return;
  double Jli[9];
  so3_left_jac_inv(lg, Jli); // argument is -logRotErr = lg
  if (J[1]) {
    double Rc[9], M[9];
    q2R(qCorr, Rc);
    mm<3, 3, 3>(Jli, Rc, M);
    double Hv[9], Hp[9];
    hat3(dvp, Hv);
    hat3(dpp, Hp);
    double* o = J[1];
    for (int i = 0; i < 54; i++) o[i] = 0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) {
        o[i * 6 + 3 + j] = -M[i * 3 + j];
        o[(3 + i) * 6 + 3 + j] = Hv[i * 3 + j];
        o[(6 + i) * 6 + 3 + j] = Hp[i * 3 + j];
      }
    for (int i = 0; i < 3; i++) o[(6 + i) * 6 + i] = -1.0;
  }
  if (J[2]) {
    double* o = J[2];
    for (int i = 0; i < 27; i++) o[i] = 0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) { o[(3 + i) * 3 + j] = Rp[i * 3 + j]; o[(6 + i) * 3 + j] = Rp[i * 3 + j] * dt; }
  }
  if (J[3]) {
    double Re[9], M[9];
    q2R(qErr, Re);
    mm<3, 3, 3>(Jli, Re, M);
    double* o = J[3];
    for (int i = 0; i < 54; i++) o[i] = 0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) { o[i * 6 + 3 + j] = M[i * 3 + j]; o[(6 + i) * 6 + j] = Rpn[i * 3 + j]; }
  }
  if (J[4]) {
    double* o = J[4];
    for (int i = 0; i < 27; i++) o[i] = 0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) o[(3 + i) * 3 + j] = -Rp[i * 3 + j];
  }
  if (J[5]) { // gravity (S2), 9x2: [0; dt Rp O^T; dt^2/2 Rp O^T]
    double O[6];
    s2_ortho(g, O);
    double* o = J[5];
    for (int i = 0; i < 18; i++) o[i] = 0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 2; j++) {
        const double ro = Rp[i * 3] * O[j * 3] + Rp[i * 3 + 1] * O[j * 3 + 1] + Rp[i * 3 + 2] * O[j * 3 + 2];
        o[(3 + i) * 2 + j] = dt * ro;
        o[(6 + i) * 2 + j] = dt * ro * (0.5 * dt);
      }
  }
  if (J[0]) {
    double Jl[9], M[9];
    so3_left_jac(nc, Jl);
    mm<3, 3, 3>(Jli, Jl, M);
    double* o = J[0];
    for (int i = 0; i < 3; i++)
      for (int k = 0; k < cdim; k++)
        o[i * cdim + k] = M[i * 3] * PJ[k] + M[i * 3 + 1] * PJ[23 + k] + M[i * 3 + 2] * PJ[46 + k];
    for (int i = 3; i < 9; i++)
      for (int k = 0; k < cdim; k++) o[i * cdim + k] = PJ[i * 23 + k];
  }
}

// r = D (.) delta padded to E rows; Jprev = -diag(D), Jnext = +diag(D)  (E x tg)
VB_DEV void diag_residual(const double* delta, const double* D, int tg, int E, double* r, double* Jp, double* Jn) {
  for (int i = 0; i < E; i++) r[i] = 0;
  for (int i = 0; i < tg; i++) r[i] = delta[i] * D[i];
  if (Jp) { for (int i = 0; i < E * tg; i++) Jp[i] = 0; for (int i = 0; i < tg; i++) Jp[i * tg + i] = -D[i]; }
  if (Jn) { for (int i = 0; i < E * tg; i++) Jn[i] = 0; for (int i = 0; i < tg; i++) Jn[i * tg + i] = D[i]; }
}
// RandomWalkFactor.cpp:107-123,141-157: r = D (.) log(next * prev^-1)
VB_DEV void rw_se3(const double* prev, const double* next, const double* D, double* r, double* const* J) {
  double pi[7], E[7], lg[6];
  se3_inv(prev, pi);
  se3_mul(next, pi, E);
  se3_log(E, lg);
  for (int i = 0; i < 6; i++) r[i] = lg[i] * D[i];
  if (!J) return;
  double Jli[36];
  se3_left_jac_inv(lg, Jli);
  if (J[0]) {
    double A[36], M[36];
    se3_adj(E, A);
    mm<6, 6, 6>(Jli, A, M);
    for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) J[0][i * 6 + j] = -D[i] * M[i * 6 + j];
  }
  if (J[1]) for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) J[1][i * 6 + j] = D[i] * Jli[i * 6 + j];
}
// PriorFactor.cpp:148-155,167-174: r = D (.) log(T * prior^-1); c = D(6), priorInv(7)
VB_DEV void prior_se3(const double* x, const double* c, double* r, double* const* J) {
  double E[7], lg[6];
  se3_mul(x, c + 6, E);
  se3_log(E, lg);
  for (int i = 0; i < 6; i++) r[i] = lg[i] * c[i];
  if (J && J[0]) {
    double Jli[36];
    se3_left_jac_inv(lg, Jli);
    for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) J[0][i * 6 + j] = c[i] * Jli[i * 6 + j];
  }
}
#define VB_OMEGA_PRIOR_STD (10.0 * 3.14159265358979323846 / 180.0) /* viba/common/Constants.h:19 */
// OmegaPriorFactor.cpp:24-31
VB_DEV void omega_prior(const double* w, const double* c, double* r, double* const* J) {
  for (int i = 0; i < 3; i++) r[i] = (w[i] - c[i]) / VB_OMEGA_PRIOR_STD;
  if (J && J[0]) { for (int i = 0; i < 9; i++) J[0][i] = 0; for (int i = 0; i < 3; i++) J[0][i * 3 + i] = 1.0 / VB_OMEGA_PRIOR_STD; }
}
// OmegaPriorFactor.cpp:35-53: r = (w - R^T w_meas)/std ; J_T = [0, -R^T hat(w_meas)/std]
VB_DEV void omega_prior_sec(const double* w, const double* T, const double* c, double* r, double* const* J) {
  double R[9], wb[3];
  q2R(T, R);
  m3tv(R, c, wb);
  for (int i = 0; i < 3; i++) r[i] = (w[i] - wb[i]) / VB_OMEGA_PRIOR_STD;
  if (!J) return;
  if (J[0]) { for (int i = 0; i < 9; i++) J[0][i] = 0; for (int i = 0; i < 3; i++) J[0][i * 3 + i] = 1.0 / VB_OMEGA_PRIOR_STD; }
  if (J[1]) {
    double H[9];
    hat3(c, H);
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) {
        const double m = R[0 * 3 + i] * H[0 * 3 + j] + R[1 * 3 + i] * H[1 * 3 + j] + R[2 * 3 + i] * H[2 * 3 + j];
        J[1][i * 6 + j] = 0;
        J[1][i * 6 + 3 + j] = -m / VB_OMEGA_PRIOR_STD;
      }
  }
}

} // namespace vb
