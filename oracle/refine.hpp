// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// refinePoints: the per-landmark 3x3 Levenberg-Marquardt pre-pass of
//   viba/problem/PointRefinement.cpp:96-196 (optimizeOnePoint, refinePoints), with
//   pointGradHess / pointCost (:49-86) over FactorStore::varGradHess (lib/small_thing/Factor.h:419-468) and
//   FactorStore::singleCost (Factor.h:230-232 -> computeSingleCost(i, false), :390-417).
// The reference walks an unordered_map<PointVariable*, FactorList> (implementation-defined order; the landmarks are
// independent, so only the order of the totals' summation depends on it) whose lists hold, store by store, the factors
// of each landmark in factor order; here: landmarks in variable order, stores in graph order.
// Eigen's 3x3 ldlt() is restated as a Cholesky solve of the damped (SPD) block.
#pragma once
#include <vector>
#include "engine.hpp"

namespace orc {

struct RefineStats {
  double totalStartCost = 0, totalEndCost = 0;
  int64_t nPoints = 0, nFailures = 0, nTotalIterations = 0, nAtLeastOneIt = 0;
};

inline bool solve3(const double* H, const double* g, double* x) {   // H x = g, H symmetric positive definite
  const double l00 = std::sqrt(H[0]);
  if (!(l00 > 0)) return false;
  const double l10 = H[3] / l00, l20 = H[6] / l00;
  const double d1 = H[4] - l10 * l10;
  if (!(d1 > 0)) return false;
  const double l11 = std::sqrt(d1), l21 = (H[7] - l20 * l10) / l11;
  const double d2 = H[8] - l20 * l20 - l21 * l21;
  if (!(d2 > 0)) return false;
  const double l22 = std::sqrt(d2);
  const double y0 = g[0] / l00, y1 = (g[1] - l10 * y0) / l11, y2 = (g[2] - l20 * y0 - l21 * y1) / l22;
  x[2] = y2 / l22;
  x[1] = (y1 - l21 * x[2]) / l11;
  x[0] = (y0 - l10 * x[1] - l20 * x[2]) / l00;
  return true;
}

// per landmark out: startCost, endCost, nIts (as double); landmark values are updated in P.data
inline RefineStats refinePoints(Engine& e, double* perPoint /* 3 x npoints or null */) {
  Problem& P = e.P;
  constexpr double kLambda = 1e-5, kCostTolerance = 1e-8, kStepTolerance = 1e-6, kMinImprovementRatio = 0.2, kStepReduction = 0.3;
  constexpr int kMaxNumIterations = 5;
  // per landmark variable: (store, factor) list, store by store, factor order
  std::vector<std::vector<std::pair<int, int64_t>>> lists(P.nv);
  for (size_t s = 0; s < P.stores.size(); s++) {
    const Store& st = P.stores[s];
    if (st.type != VB_F_VISUAL && st.type != VB_F_VISUAL_RS) continue;
    for (int64_t k = 0; k < st.count; k++) lists[st.vars[k * st.nvars]].push_back({(int)s, k});
  }
  RefineStats out;
  std::vector<double> backups;
  int64_t pi = 0;
  for (int64_t v = 0; v < P.nv; v++) {
    const auto& fl = lists[v];
    if (fl.empty()) continue;
    double* X = P.data.data() + P.off[v];
    auto gradHess = [&](bool update, double* grad, double* hess) {
      double cost = 0;
      for (int i = 0; i < 3; i++) grad[i] = 0;
      for (int i = 0; i < 9; i++) hess[i] = 0;
      backups.resize(fl.size());
      for (size_t q = 0; q < fl.size(); q++) {
        const Store& st = P.stores[fl[q].first];
        double r[2], JX[6];
        double* J[kMaxVars] = {JX, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
        if (!e.evalFactor(st, fl[q].second, r, J)) { if (update) backups[q] = -1.0; continue; }
        double soft, d;
        Engine::lossJet2(st, r[0] * r[0] + r[1] * r[1], soft, d);
        for (int a = 0; a < 3; a++) {
          grad[a] += d * (r[0] * JX[a] + r[1] * JX[3 + a]);
          for (int b = 0; b < 3; b++) hess[a * 3 + b] += d * (JX[a] * JX[b] + JX[3 + a] * JX[3 + b]);
        }
        cost += 0.5 * soft;
        if (update) backups[q] = 0.5 * soft;
      }
      return cost;
    };
    auto pointCost = [&]() {
      double c = 0;
      for (size_t q = 0; q < fl.size(); q++) {
        if (backups[q] < 0) continue;
        const Store& st = P.stores[fl[q].first];
        double r[2];
        if (!e.evalFactor(st, fl[q].second, r, nullptr)) { c += backups[q]; continue; }
        double soft, d;
        Engine::lossJet2(st, r[0] * r[0] + r[1] * r[1], soft, d);
        c += 0.5 * soft;
      }
      return c;
    };
    double startCost = 0, endCost = 0;
    int nIts = 0;
    for (int it = 0; it < kMaxNumIterations; it++) {
      double grad[3], hess[9];
      const double cost = gradHess(true, grad, hess);
      if (it == 0) startCost = endCost = cost;
      for (int a = 0; a < 3; a++) hess[a * 4] = hess[a * 4] * (1.0 + kLambda) + kLambda;
      double step[3];
      if (!solve3(hess, grad, step)) { nIts = -1; break; }
      for (int a = 0; a < 3; a++) step[a] = -step[a];
      const double bak[3] = {X[0], X[1], X[2]};
      bool success = false;
      for (int a = 0; a < 3; a++) X[a] += step[a];
      double newCost = pointCost();
      const double modelCostDelta = step[0] * grad[0] + step[1] * grad[1] + step[2] * grad[2];
      if (-modelCostDelta < kCostTolerance) break;   // nothing to do (the step stays applied, as in the reference)
      if (newCost < cost + modelCostDelta * kMinImprovementRatio) {
        success = true;
      } else {
        double nGrad[3], nHess[9];
        gradHess(false, nGrad, nHess);
        for (int a = 0; a < 3; a++) X[a] = bak[a];
        const double newDelta = step[0] * nGrad[0] + step[1] * nGrad[1] + step[2] * nGrad[2];
        const double f = newDelta > 0 ? -modelCostDelta / (newDelta - modelCostDelta) : kStepReduction;
        for (int a = 0; a < 3; a++) step[a] *= f;
        for (int a = 0; a < 3; a++) X[a] += step[a];
        newCost = pointCost();
        if (newCost < cost + (step[0] * grad[0] + step[1] * grad[1] + step[2] * grad[2]) * kMinImprovementRatio) success = true;
        else for (int a = 0; a < 3; a++) X[a] = bak[a];
      }
      if (success) { nIts++; endCost = newCost; }
      else { nIts = -1; break; }
      if (step[0] * step[0] + step[1] * step[1] + step[2] * step[2] < kStepTolerance * kStepTolerance) break;
    }
    if (nIts < 0) out.nFailures++;
    else { out.nTotalIterations += nIts; out.nAtLeastOneIt += nIts > 0; }
    out.totalStartCost += startCost;
    out.totalEndCost += endCost;
    out.nPoints++;
    if (perPoint) { perPoint[3 * pi] = startCost; perPoint[3 * pi + 1] = endCost; perPoint[3 * pi + 2] = nIts; }
    pi++;
  }
  return out;
}

// refineTriangulationResult (viba/single_session/Triangulation.cpp:99-161), the refinement half of triangulatePoint
// (:165-240), applied to every landmark over its visual observations: maxNumIterations undamped Gauss-Newton steps on the
// point with small_thing::HuberLoss(radius) (SoftLoss.h:64-113) on the weighted reprojection error; every iteration
// re-classifies the observations — inlier = squared UNWEIGHTED image error below outlierThreshold^2 (:139-144) — and with
// skipOutliers the outliers do not enter the normal equations.  The residual and d(residual)/d(point) are the visual
// factors' own (VisualFactor / RollingShutterVisualFactor use the observed image row for the rolling-shutter pose,
// VisualFactor.cpp:142-150, exactly as T_bodyImu_world_atImageRow does, :303-327); the unweighted error is recovered from
// the weighted one through sqrtH^-1.  numInlierObservations / the inlier list are those of the LAST iteration, i.e. at
// the point before the last update, as in the reference.  Eigen's llt() of a non-SPD 3x3 returns garbage there; here such
// a landmark is left where it is and reported with status -1.
// perPoint: {numInliers, cost of the last iteration, status}; inlier[s][k] = 1 for the inlier factors of store s.
inline int64_t refineTriangulation(Engine& e, double outlierThreshold, bool skipOutliers, int maxNumIterations, double huberRadius,
                                   double* perPoint /* 3 x npoints or null */, std::vector<std::vector<uint8_t>>* inlier) {
  Problem& P = e.P;
  const double thr2 = outlierThreshold * outlierThreshold, b = huberRadius * huberRadius;
  std::vector<std::vector<std::pair<int, int64_t>>> lists(P.nv);
  if (inlier) inlier->assign(P.stores.size(), {});
  for (size_t s = 0; s < P.stores.size(); s++) {
    const Store& st = P.stores[s];
    if (st.type != VB_F_VISUAL && st.type != VB_F_VISUAL_RS) continue;
    if (inlier) (*inlier)[s].assign((size_t)st.count, 0);
    for (int64_t k = 0; k < st.count; k++) lists[st.vars[k * st.nvars]].push_back({(int)s, k});
  }
  int64_t pi = 0, failures = 0;
  for (int64_t v = 0; v < P.nv; v++) {
    const auto& fl = lists[v];
    if (fl.empty()) continue;
    double* X = P.data.data() + P.off[v];
    double nIn = 0, cost = 0;
    int status = 0;
    for (int it = 0; it < maxNumIterations; it++) {
      double grad[3] = {0, 0, 0}, hess[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      cost = 0; nIn = 0;
      for (size_t q = 0; q < fl.size(); q++) {
        const Store& st = P.stores[fl[q].first];
        const int64_t k = fl[q].second;
        if (inlier) (*inlier)[fl[q].first][(size_t)k] = 0;
        double r[2], JX[6];
        double* J[kMaxVars] = {JX, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
        if (!e.evalFactor(st, k, r, J)) continue;
        const double* c = st.consts.data() + k * (int64_t)st.cpf;   // uv(2), sqrtH(4) row-major, ...
        const double det = c[2] * c[5] - c[3] * c[4];
        const double e0 = (c[5] * r[0] - c[3] * r[1]) / det, e1 = (-c[4] * r[0] + c[2] * r[1]) / det;
        if (e0 * e0 + e1 * e1 < thr2) { nIn += 1; if (inlier) (*inlier)[fl[q].first][(size_t)k] = 1; }
        else if (skipOutliers) continue;
        const double s2 = r[0] * r[0] + r[1] * r[1];
        double soft = s2, d = 1.0;
        if (s2 > b) { const double rr = std::sqrt(s2); soft = 2.0 * huberRadius * rr - b; d = huberRadius / rr; }
        cost += 0.5 * soft;
        for (int a = 0; a < 3; a++) {
          grad[a] += d * (r[0] * JX[a] + r[1] * JX[3 + a]);
          for (int bb = 0; bb < 3; bb++) hess[a * 3 + bb] += d * (JX[a] * JX[bb] + JX[3 + a] * JX[3 + bb]);
        }
      }
      double step[3];
      if (!solve3(hess, grad, step)) { status = -1; break; }
      for (int a = 0; a < 3; a++) X[a] -= step[a];
    }
    failures += status < 0;
    if (perPoint) { perPoint[3 * pi] = nIn; perPoint[3 * pi + 1] = cost; perPoint[3 * pi + 2] = status; }
    pi++;
  }
  return failures;
}

}  // namespace orc
