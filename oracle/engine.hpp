// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// Single-threaded (deterministic, numThreads<=1 ordering: Factor.h:726-733) restatement of
//   FactorStore::computeSingleGradHess / computeSingleCost   lib/small_thing/Factor.h:390-417,543-661
//   Optimizer::initSolver block set -> sorted lower block CSR  lib/small_thing/Optimizer.cpp:166-207
//   Optimizer::addDamping / applyStep / backup / restore       lib/small_thing/Optimizer.cpp:107-146
// The Hessian lives in a lower block-CSR keyed by registration index (row i, columns j <= i,
// diagonal last in each row exactly as Optimizer.cpp:186-197 builds it); each block is row-major
// d_i x d_j.  grad/step are in registration order (the oracle applies no permutation; the
// product's permuted vectors are compared through vb_get_param_start).
#pragma once
#include <set>
#include <utility>
#include <string>
#include <stdexcept>
#include <functional>
#include "factors.hpp"
#include "factors_small.hpp"
#include "factors_ext.hpp"

namespace orc {

struct CostStats { int64_t numTotal = 0, numInvalid = 0, numPrevInvalid = 0; };

struct Engine {
  Problem P;
  // lower block CSR
  std::vector<int64_t> rowPtr, colInd, blkOff; // blkOff[k] offset in hess
  std::vector<int64_t> paramStart;            // scalar offset of each registration index
  int64_t order = 0, dataSize = 0;
  std::vector<double> hess, grad, step, gradNew;

  bool optionalErrors(int type) const { return type == VB_F_VISUAL || type == VB_F_VISUAL_RS; }

  // ---- functor dispatch -------------------------------------------------------------------
  // returns false for an invalid (optional-empty) residual
  bool evalFactor(const Store& st, int64_t k, double* r, double** J) const {
    const int32_t* v = &st.vars[k * st.nvars];
    const double* c = st.consts.data() + k * (int64_t)st.cpf;
    auto D = [&](int a) { return P.data.data() + P.off[v[a]]; };
    switch (st.type) {
      case VB_F_VISUAL:
        return visualFactor(D(0), D(1), D(2), D(3), c, P.tdim[v[3]], r, J);
      case VB_F_VISUAL_RS:
        return rsVisualFactor(D(0), D(1), D(2), D(3), P.flags[v[3]], P.tdim[v[3]], D(4), c, st.aux.data(), r, J);
      case VB_F_INERTIAL_SEC:
        secondaryInertialFactorCommon(D(0), P.flags[v[0]], P.tdim[v[0]], D(1), D(2), D(3), D(4), D(5), D(6), D(7), D(8),
                                      st.aux.data() + k * kPreintStride, r, J);
        return true;
      case VB_F_INERTIAL_SEC_SPLIT:
        secondaryInertialFactor(D(0), P.flags[v[0]], P.tdim[v[0]], D(1), D(2), D(3), D(4), D(5), D(6), D(7), D(8), D(9),
                                st.aux.data() + k * kPreintStride, r, J);
        return true;
      case VB_F_INERTIAL:
        inertialFactor(D(0), P.flags[v[0]], P.tdim[v[0]], D(1), D(2), D(3), D(4), D(5),
                       st.aux.data() + k * kPreintStride, r, J);
        return true;
      case VB_F_RW_IMUCALIB: rwImuCalib(D(0), D(1), P.flags[v[0]], P.tdim[v[0]], c, r, J); return true;
      case VB_F_RW_CAM: rwCam(D(0), D(1), P.flags[v[0]], P.tdim[v[0]], c, r, J); return true;
      case VB_F_RW_SE3: rwSE3(D(0), D(1), c, r, J); return true;
      case VB_F_PRIOR_IMUCALIB: priorImuCalib(D(0), P.flags[v[0]], P.tdim[v[0]], c, r, J); return true;
      case VB_F_PRIOR_CAM: priorCam(D(0), P.flags[v[0]], P.tdim[v[0]], c, r, J); return true;
      case VB_F_PRIOR_SE3: priorSE3(D(0), c, r, J); return true;
      case VB_F_OMEGA_PRIOR: omegaPrior(D(0), c, r, J); return true;
      case VB_F_OMEGA_PRIOR_SEC: omegaPriorSec(D(0), D(1), c, r, J); return true;
    }
    throw std::runtime_error("oracle: unsupported factor type " + std::to_string(st.type));
  }

  // HuberLossWithCutoff — lib/small_thing/SoftLoss.h:115-163 ; a<=0 => TrivialLoss
  static void lossJet2(const Store& st, double s, double& val, double& der) {
    if (!(st.loss_a > 0) || std::isinf(st.loss_a)) { val = s; der = 1.0; return; }
    const double a = st.loss_a, b = a * a, k2 = st.loss_k * st.loss_k, h = 2.0 * a * st.loss_k - b;
    if (s > k2) { val = h; der = 0.0; }
    else if (s > b) { const double r = std::sqrt(s); val = 2.0 * a * r - b; der = a / r; }
    else { val = s; der = 1.0; }
  }
  // squaredError: r^T Lambda r (Factor.h:143-152)
  static double squaredError(const Store& st, int64_t k, const double* r, int E) {
    if (!factorHasPrecision(st.type)) { double s = 0; for (int i = 0; i < E; i++) s += r[i] * r[i]; return s; }
    const double* L = st.consts.data() + k * (int64_t)st.cpf; // 9x9
    double s = 0;
    for (int i = 0; i < E; i++) { double t = 0; for (int j = 0; j < E; j++) t += L[i * E + j] * r[j]; s += r[i] * t; }
    return s;
  }

  // ---- symbolic: Optimizer::initSolver (Optimizer.cpp:166-207) ------------------------------
  void symbolic() {
    std::set<std::pair<int64_t, int64_t>> blocks;
    for (const Store& st : P.stores)
      for (int64_t k = 0; k < st.count; k++) {
        const int32_t* v = &st.vars[k * st.nvars];
        for (int i = 0; i < st.nvars; i++) {
          const int64_t ii = P.index[v[i]];
          if (ii < 0) continue;
          for (int j = 0; j < i; j++) {
            const int64_t jj = P.index[v[j]];
            if (jj < 0 || jj == ii) continue;
            blocks.insert({std::max(ii, jj), std::min(ii, jj)});
          }
        }
      }
    const int64_t n = P.nblocks;
    rowPtr.assign(n + 1, 0);
    colInd.clear();
    auto it = blocks.begin();
    for (int64_t i = 0; i < n; i++) {
      rowPtr[i] = (int64_t)colInd.size();
      while (it != blocks.end() && it->first == i) { colInd.push_back(it->second); ++it; }
      colInd.push_back(i); // explicit diagonal, appended after the off-diagonals
    }
    rowPtr[n] = (int64_t)colInd.size();
    paramStart.assign(n + 1, 0);
    for (int64_t i = 0; i < n; i++) paramStart[i + 1] = paramStart[i] + P.blockSize[i];
    order = paramStart[n];
    blkOff.assign(colInd.size(), 0);
    int64_t off = 0;
    for (int64_t i = 0; i < n; i++)
      for (int64_t k = rowPtr[i]; k < rowPtr[i + 1]; k++) { blkOff[k] = off; off += (int64_t)P.blockSize[i] * P.blockSize[colInd[k]]; }
    dataSize = off;
    hess.assign(dataSize, 0.0);
    grad.assign(order, 0.0);
    step.assign(order, 0.0);
    gradNew.assign(order, 0.0);
  }
  // block (i,j), i >= j: offset in hess or -1
  int64_t findBlock(int64_t i, int64_t j) const {
    const int64_t* b = &colInd[rowPtr[i]];
    const int64_t* e = &colInd[rowPtr[i + 1]];
    const int64_t* p = std::lower_bound(b, e, j);
    if (p == e || *p != j) return -1;
    return blkOff[p - colInd.data()];
  }

  // CPU-baseline mode (bench.py only): factor loop over `threads` OpenMP threads with atomic block
  // adds, the way the reference runs it on its thread pool with locked adds (Factor.h:704-734,
  // AtomicOps.h:80-112).  threads == 1 is the serial checker path the tests use (bit-stable).
  int threads = 1;
  std::function<void()> preStep; // Settings::preStepCallback (Optimizer.cpp:801-805): data refresh before every iteration
  inline void acc(double& dst, double v) const {
    if (threads > 1) {
#pragma omp atomic
      dst += v;
    } else dst += v;
  }
  // ---- computeSingleGradHess (Factor.h:543-661) ------------------------------------------------
  double singleGradHess(const Store& st, int64_t k, double* g, double* H, bool updateCache, bool dontRetryFailed) {
    const bool opt = optionalErrors(st.type);
    if (opt && dontRetryFailed && st.cache[k] < 0.0) return 0.0;
    const int E = factorErrorSize(st.type);
    const int32_t* v = &st.vars[k * st.nvars];
    double r[kMaxE];
    double Jbuf[kMaxVars][kMaxE * 23];
    double* J[kMaxVars];
    for (int i = 0; i < st.nvars; i++) J[i] = P.index[v[i]] < 0 ? nullptr : Jbuf[i];
    if (!evalFactor(st, k, r, J)) {
      if (updateCache || dontRetryFailed) st.cache[k] = -1.0;
      return 0.0;
    }
    double soft, dSoft;
    lossJet2(st, squaredError(st, k, r, E), soft, dSoft);
    const double* L = factorHasPrecision(st.type) ? st.consts.data() + k * (int64_t)st.cpf : nullptr;
    double adj[kMaxE * 23], blk[23 * 23];
    for (int i = 0; i < st.nvars; i++) {
      const int64_t ii = P.index[v[i]];
      if (ii < 0) continue;
      const int di = P.tdim[v[i]];
      if (L) matmul(L, J[i], adj, E, E, di); else std::memcpy(adj, J[i], sizeof(double) * E * di);
      for (int t = 0; t < E * di; t++) adj[t] *= dSoft;
      double* gs = g + paramStart[ii];
      for (int a = 0; a < di; a++) { double s = 0; for (int e = 0; e < E; e++) s += r[e] * adj[e * di + a]; acc(gs[a], s); }
      if (!H) continue;
      for (int j = 0; j < i; j++) {
        const int64_t jj = P.index[v[j]];
        if (jj < 0) continue;
        const int dj = P.tdim[v[j]];
        matmulTN(adj, J[j], blk, E, di, dj); // iAdjJac^T * jJac  (di x dj)
        if (ii > jj) {
          double* B = H + findBlock(ii, jj);
          for (int t = 0; t < di * dj; t++) acc(B[t], blk[t]);
        } else if (ii < jj) { // stored transposed as block (jj, ii): dj x di
          double* B = H + findBlock(jj, ii);
          for (int a = 0; a < di; a++) for (int b = 0; b < dj; b++) acc(B[b * di + a], blk[a * dj + b]);
        } else { // same variable twice: diagonal block gets the off-diagonal product too
          double* B = H + findBlock(ii, ii);
          for (int t = 0; t < di * dj; t++) acc(B[t], blk[t]);
        }
      }
      matmulTN(adj, J[i], blk, E, di, di);
      double* B = H + findBlock(ii, ii);
      for (int t = 0; t < di * di; t++) acc(B[t], blk[t]);
    }
    const double ret = soft * 0.5;
    if (opt && updateCache) st.cache[k] = ret;
    return ret;
  }

  // Optimizer::computeGradHess (Optimizer.cpp:57-71)
  double computeGradHess(double* g, double* H, bool updateCache, bool dontRetryFailed) {
    double tot = 0;
    for (const Store& st : P.stores) {
      double c = 0;
      if (threads > 1) {
#pragma omp parallel for schedule(dynamic, 512) reduction(+ : c) num_threads(threads)
        for (int64_t k = 0; k < st.count; k++) c += singleGradHess(st, k, g, H, updateCache, dontRetryFailed);
      } else {
        for (int64_t k = 0; k < st.count; k++) c += singleGradHess(st, k, g, H, updateCache, dontRetryFailed);
      }
      tot += c;
    }
    return tot;
  }

  // computeSingleCost (Factor.h:390-417); returns has_value via `has`
  double singleCost(const Store& st, int64_t k, bool comparable, CostStats* stats, bool& has) const {
    const int E = factorErrorSize(st.type);
    double r[kMaxE];
    const bool valid = evalFactor(st, k, r, nullptr);
    if (stats) stats->numTotal += 1;
    has = true;
    if (optionalErrors(st.type)) {
      const double prev = st.cache[k];
      const bool prevInvalid = prev < 0.0;
      if (stats) { stats->numInvalid += valid ? 0 : 1; stats->numPrevInvalid += prevInvalid ? 1 : 0; }
      if (comparable) {
        if (prevInvalid) { has = false; return 0.0; }
        if (!valid) return prev;
      }
    }
    if (!valid) { has = false; return 0.0; }
    double val, der;
    lossJet2(st, squaredError(st, k, r, E), val, der);
    return val * 0.5;
  }
  // Optimizer::computeCost (Optimizer.cpp:88-97); NOTE Q3: stats are overwritten per store
  double computeCost(bool comparable, CostStats* stats) const {
    double tot = 0;
    for (const Store& st : P.stores) {
      CostStats cs;
      double c = 0;
      if (threads > 1) {
        int64_t nt = 0, ni = 0, np = 0;
#pragma omp parallel for schedule(static) reduction(+ : c, nt, ni, np) num_threads(threads)
        for (int64_t k = 0; k < st.count; k++) {
          bool has; CostStats one;
          c += singleCost(st, k, comparable, &one, has);
          nt += one.numTotal; ni += one.numInvalid; np += one.numPrevInvalid;
        }
        cs.numTotal = nt; cs.numInvalid = ni; cs.numPrevInvalid = np;
      } else
      for (int64_t k = 0; k < st.count; k++) { bool has; c += singleCost(st, k, comparable, &cs, has); }
      if (stats) *stats = cs;
      tot += c;
    }
    return tot;
  }

  // Optimizer::addDamping (Optimizer.cpp:136-146)
  void addDamping(double lambda) {
    for (int64_t i = 0; i < P.nblocks; i++) {
      double* B = hess.data() + blkOff[rowPtr[i + 1] - 1];
      const int d = P.blockSize[i];
      for (int a = 0; a < d; a++) B[a * d + a] = B[a * d + a] * (1.0 + lambda) + lambda;
    }
  }
  // Optimizer::applyStep (Optimizer.cpp:121-134)
  void applyStep(const double* s, double ratios[3]) {
    double mx = 0, sq = 0, sm = 0;
    int64_t n = 0;
    for (int64_t b = 0; b < P.nblocks; b++) {
      const int64_t v = P.blockVar[b];
      if (v < 0) continue;
      const double r = applyBoxPlus(P.kind[v], P.flags[v], P.aux[v], P.data.data() + P.off[v], s + paramStart[b], P.tdim[v]);
      mx = std::max(mx, r); sq += r * r; sm += r; n++;
    }
    ratios[0] = mx; ratios[1] = std::sqrt(sq / n); ratios[2] = sm / n;
  }
  void backup() { P.backup = P.data; }
  void restore() { P.data = P.backup; }
};

} // namespace orc
