// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// Optimizer::optimize — lib/small_thing/Optimizer.cpp:768-1106, restated with the quirks of
// SURVEY.md App. A kept (Q1 fall-through on negative model reduction, Q2 stale newCost in the
// step-factor loop, Q3 stats of the last store, Q5 dontRetryFailed latch).
#pragma once
#include <cstdio>
#include <cmath>
#include <functional>
#include "solver.hpp"

namespace orc {

struct LmTrace { // one entry per iteration, for parity tests
  std::vector<double> prevCost, newCost, damping, modelRed;
  std::vector<vb_iter_trace> full; // the fields of the reference's log line + the branches taken
};

inline double vdot(const std::vector<double>& a, const std::vector<double>& b) {
  double s = 0;
  for (size_t i = 0; i < a.size(); i++) s += a[i] * b[i];
  return s;
}

inline int optimize(Engine& e, SchurSkylineSolver& sol, const vb_settings& st, vb_summary* out, LmTrace* trace,
                    const std::function<void(int)>& userPreStep = nullptr) {
  double damping = st.damping;
  int iterationNum = 0, lastImprovementIteration = 0, lastTroubledIteration = -10;
  double initialCost = 0, finalCost = 0, troubledSeqStartDamping = damping;
  int troubledSeqStart = 0, numTroubledSeqs = 0, largestTroubledSeq = 0, stopReason = 1;
  bool dontRetryFailed = false;
  std::vector<double> hessF; // factor() works on a copy: the oracle solver is not in-place
  auto acceptable = [](const CostStats& s) {
    const double rate = s.numInvalid / (s.numTotal + 1.0);
    return rate < 0.03 && (s.numInvalid < s.numPrevInvalid * 2.0 + 50);
  };
  sol.setLinearSolver(st.solver_type, st.pcg_max_iterations, st.pcg_desired_residual);   // Settings::solverType (Optimizer.h:43)
  while (true) {
    int branch = 0, attempts = 0;
    if (userPreStep) userPreStep(iterationNum); // Settings::preStepCallback (Optimizer.cpp:803-805)
    if (e.preStep) e.preStep();
    const double dampingUsed = damping;
    std::fill(e.grad.begin(), e.grad.end(), 0.0);
    std::fill(e.hess.begin(), e.hess.end(), 0.0);
    const double prevCost = e.computeGradHess(e.grad.data(), e.hess.data(), true, dontRetryFailed);
    finalCost = prevCost;
    if (iterationNum == 0) initialCost = prevCost;
    e.addDamping(damping);
    hessF = e.hess;
    sol.factorAndSolve(hessF.data(), e.grad.data(), e.step.data());
    double modelCostReduction = vdot(e.step, e.grad) * 0.5;
    if (modelCostReduction < 0) { // Q1
      std::fill(e.grad.begin(), e.grad.end(), 0.0);
      std::fill(e.hess.begin(), e.hess.end(), 0.0);
      e.computeGradHess(e.grad.data(), e.hess.data(), true, dontRetryFailed);
      damping *= st.damping_adjust_on_fail;
      branch |= VB_BR_MODEL_NEGATIVE;
    }
    for (double& s : e.step) s = -s;
    const double gradNorm = std::sqrt(vdot(e.grad, e.grad)), stepNorm = std::sqrt(vdot(e.step, e.step));
    e.backup();
    double ratios[3];
    e.applyStep(e.step.data(), ratios);
    CostStats stats;
    double newCost = e.computeCost(true, &stats);
    double costReduction = prevCost - newCost;
    const double ratioReductionToCost = costReduction / newCost;
    double ratioReductionToExpected = costReduction / modelCostReduction;
    double appliedStepFactor = 1.0;
    bool failureRateIsAcceptable = acceptable(stats);
    if (st.max_step_factor_attempts > 0 &&
        (ratioReductionToExpected < st.min_relative_cost_reduction || !failureRateIsAcceptable)) {
      branch |= VB_BR_STEP_FACTOR;
      std::fill(e.gradNew.begin(), e.gradNew.end(), 0.0);
      e.computeGradHess(e.gradNew.data(), nullptr, false, dontRetryFailed);
      const double backRed = -vdot(e.gradNew, e.step) * 0.5;
      double stepFactor = backRed > 0 ? (modelCostReduction / (modelCostReduction + backRed)) : st.step_factor_decrease;
      for (int i = 0; i < st.max_step_factor_attempts; i++) {
        attempts++;
        appliedStepFactor *= stepFactor;
        for (double& s : e.step) s *= stepFactor;
        e.restore();
        double rr[3];
        e.applyStep(e.step.data(), rr);
        CostStats statsAtF;
        const double newCostAtF = e.computeCost(true, &statsAtF);
        const double costReductionAtF = prevCost - newCost; // Q2
        const double ratioAtF = costReductionAtF / (modelCostReduction * appliedStepFactor);
        if (ratioAtF >= st.min_relative_cost_reduction && acceptable(statsAtF)) {
          newCost = newCostAtF; stats = statsAtF; costReduction = costReductionAtF;
          ratioReductionToExpected = ratioAtF; failureRateIsAcceptable = true;
          branch |= VB_BR_ACCEPT_AT_FACTOR;
          break;
        }
        if (st.try_sub_step) {
          branch |= VB_BR_SUB_STEP;
          std::fill(e.gradNew.begin(), e.gradNew.end(), 0.0);
          e.computeGradHess(e.gradNew.data(), nullptr, false, dontRetryFailed);
          std::vector<double> g2 = e.gradNew;
          sol.solveOnly(hessF.data(), g2.data(), e.gradNew.data());
          for (double& s : e.gradNew) s = -s;
          e.applyStep(e.gradNew.data(), rr);
          CostStats statsAtSub;
          const double newCostAtSub = e.computeCost(true, &statsAtSub);
          const double costReductionAtSub = prevCost - newCostAtSub;
          const double ratioAtSub = costReductionAtSub / (modelCostReduction * appliedStepFactor);
          if (ratioAtSub >= st.min_relative_cost_reduction && acceptable(statsAtSub)) {
            newCost = newCostAtSub; stats = statsAtSub; costReduction = costReductionAtSub;
            ratioReductionToExpected = ratioAtSub; failureRateIsAcceptable = true;
            branch |= VB_BR_ACCEPT_AT_SUB;
            break;
          }
        }
        if (!dontRetryFailed) { dontRetryFailed = true; branch |= VB_BR_LATCH_NO_RETRY; }
        stepFactor = st.step_factor_decrease;
      }
    }
    const bool toleranceHit = ratioReductionToCost < st.relative_cost_tolerance ||
        costReduction < st.absolute_cost_tolerance || ratios[1] < st.variables_tolerance;
    if (toleranceHit) branch |= VB_BR_TOLERANCE_HIT;
    const double dampingLogged0 = damping;
    if (st.verbose)
      std::printf("[oracle] it %d cost %.12g -> %.12g lambda %.3g model %.6g\n", iterationNum, prevCost, newCost, damping, modelCostReduction);
    bool stopDamping = false;
    if (newCost > prevCost || !failureRateIsAcceptable) {
      if (lastTroubledIteration != iterationNum - 1) { troubledSeqStartDamping = damping; troubledSeqStart = iterationNum; }
      damping *= st.damping_adjust_on_fail;
      branch |= VB_BR_REJECTED;
      e.restore();
      if (damping > st.damping_max) stopDamping = true;
      else lastTroubledIteration = iterationNum;
    } else {
      if (lastTroubledIteration == iterationNum - 1) {
        if (troubledSeqStartDamping < 1e1 && damping > 1e-3) {
          numTroubledSeqs++;
          largestTroubledSeq = std::max(largestTroubledSeq, iterationNum - troubledSeqStart);
          branch |= VB_BR_TROUBLED_SEQ_END;
        }
      }
      if (ratioReductionToExpected >= st.min_relative_cost_reduction && appliedStepFactor > st.min_step_factor_for_good) {
        damping *= st.damping_adjust_on_good;
        damping = std::max(damping, st.damping_min);
        branch |= VB_BR_GOOD;
      } else {
        damping *= st.damping_adjust_on_average;
        branch |= VB_BR_AVERAGE;
      }
      finalCost = newCost;
    }
    if (trace) {
      trace->prevCost.push_back(prevCost); trace->newCost.push_back(newCost);
      trace->damping.push_back(dampingLogged0); trace->modelRed.push_back(modelCostReduction);
      vb_iter_trace t{};
      t.prev_cost = prevCost; t.new_cost = newCost; t.damping = damping; t.damping_used = dampingUsed;
      t.model_cost_reduction = modelCostReduction; t.applied_step_factor = appliedStepFactor;
      t.ratio_reduction_to_expected = ratioReductionToExpected; t.cost_reduction = costReduction;
      t.grad_norm = gradNorm; t.step_norm = stepNorm;
      t.ratio_s2v_inf = ratios[0]; t.ratio_s2v_2 = ratios[1]; t.ratio_s2v_1 = ratios[2];
      t.num_total = stats.numTotal; t.num_invalid = stats.numInvalid; t.num_prev_invalid = stats.numPrevInvalid;
      t.branch = branch; t.step_factor_attempts = attempts;
      trace->full.push_back(t);
    }
    if (stopDamping) { stopReason = 2; break; }
    iterationNum++;
    if (!toleranceHit) lastImprovementIteration = iterationNum;
    if (iterationNum >= lastImprovementIteration + st.stop_if_no_improvement_for &&
        iterationNum >= lastTroubledIteration + st.distance_from_troubled_iteration) { stopReason = 0; break; }
    else if (iterationNum >= st.max_num_iterations) { stopReason = 1; break; }
  }
  if (out) {
    out->initial_cost = initialCost; out->final_cost = finalCost;
    out->num_troubled_seqs = numTroubledSeqs; out->largest_troubled_seq = largestTroubledSeq;
    out->num_iterations = iterationNum; out->stop_reason = stopReason;
  }
  return 0;
}

} // namespace orc
