// small_thing/Optimizer.h — C++17 host mirror of small_thing::Optimizer above the C ABI of
// viba_b200.h (BASELINE north_star: "host code stays C++17 and calls CUDA through a
// thin C-ABI layer").  Same names, argument meaning and error behaviour as
// lib/small_thing/Optimizer.h:25-345 for the part of the class ark_vi_ba drives:
//
//   addFactor(F, [precisionMatrix], [loss], variables...)      Optimizer.h:115-175
//   registerVariable / registerAllVariables / unregisterAllVariables   :193-208, Optimizer.cpp:25-55
//   registeredVariablesToEliminationRange                      :211, Optimizer.cpp:41-47
//   totalVarDimensionality / totalErrorDimensionality          :213-215
//   computeCost(makeComparableWithStored, stats)               :232-235
//   optimize(settings) -> Summary                              :328-330, Optimizer.cpp:768-1106
//   Settings / Summary field names                             :40-99
//
// Ownership is the reference's: the optimizer stores raw pointers to user-owned Variables and to the
// loss; factor payloads keep pointers to user data (preintegration records, rolling-shutter tables)
// that the user may refresh between calls.  The device is the source of truth only inside a call:
// every compute entry point uploads the current host values first and optimize() writes the result
// back into the Variables before it returns.  Errors: the C ABI's status codes become
// std::runtime_error carrying vb_last_error(), like BASPACHO_CHECK / XR_CHECK in the reference.
// There is no host implementation of any of the numerics: without a usable sm_100 device the first
// compute call throws.
//
// Settings::preStepCallback (Optimizer.h:90, called at Optimizer.cpp:803-805) runs on the calling thread at the top of
// every iteration: the values of the IMU-calibration, camera-model and gravity variables (what the reference's callbacks
// read: InitCalibration.cpp:316-325, InertialFactors.cpp:19-43) are pulled from the device into the user's Variables
// first, and whatever the callback changed in the user-owned records (preintegration records, rolling-shutter tables,
// factor constants) is pushed to the device after it.  Settings::log receives the reference's per-iteration text
// (Optimizer.cpp:1058-1076).  Settings::solverType: all five types run on the device (PCG on the reduced camera system,
// Optimizer.cpp:232-331; for the Gauss-Seidel and lower-precision preconditioners see viba_b200.h, vb_settings).
//
// Differences, all forced by the device path: (1) factor types come from the catalogue of
// small_thing/Factor.h, not arbitrary functors; (2) store order is first-use order (the reference's is
// libstdc++ unordered_map order over type_index); (3) Settings::numThreads is ignored.
#pragma once

#include <cstdint>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <typeindex>
#include <unordered_map>
#include <utility>
#include <vector>

#include "../viba_b200.h"
#include "Factor.h"
#include "SoftLoss.h"
#include "Variable.h"

namespace small_thing {

using LogFunc = std::function<void(const std::string&)>;
using PreStepCallback = std::function<void(int)>;   // Optimizer.h:24

struct CostStats {   // Factor.h:25-31
  int64_t numTotal = 0, numInvalid = 0, numPrevInvalid = 0;
};

// the graph in the C ABI's terms; arrays owned here, `desc` points into them
struct ExportedGraph {
  std::vector<int32_t> kind, tdim;
  std::vector<uint32_t> flags;
  std::vector<int64_t> index, off;
  std::vector<double> aux, data;
  std::vector<VarBase*> pool;                      // pool id -> user variable
  std::vector<std::vector<int32_t>> storeVars;
  std::vector<std::vector<double>> storeConsts, storeAux;
  std::vector<vb_factor_store> stores;
  vb_graph_desc desc{};
};

class Optimizer {
 public:
  enum SolverType {   // Optimizer.h:30-36
    Solver_Direct,
    Solver_PCG_Trivial,
    Solver_PCG_Jacobi,
    Solver_PCG_GaussSeidel,
    Solver_PCG_LowerPrecSolve,
  };

  struct Settings {   // Optimizer.h:40-91, same defaults
    int maxNumIterations = 50;
    unsigned int numThreads = 8;   // ignored: one host thread drives one GPU context
    SolverType solverType = Solver_Direct;
    int pcgMaxIterations = 40;
    double pcgDesiredResidual = 1e-10;
    double absoluteCostTolerance = 1e-8;
    double relativeCostTolerance = 1e-10;
    double variablesTolerance = 1e-5;
    int stopIfNoImprovementFor = 3;
    int distanceFromTroubledIteration = 3;
    double damping = 1e-5;
    double dampingAdjustOnFail = 2.5;
    double dampingAdjustOnGoodStep = 0.7;
    double dampingAdjustOnAverageStep = 1.5;
    double dampingMax = 1e8;
    double dampingMin = 1e-9;
    double minRelativeCostReduction = 0.3;
    double stepFactorDecrease = 0.3;
    int maxStepFactorAttempts = 2;
    bool trySubStep = true;
    double minStepFactorForGood = 0.7;
    LogFunc log = [](const std::string& str) { std::cout << str << std::endl; };
    PreStepCallback preStepCallback;   // Optimizer.h:90
    bool verbose = false;          // per-iteration lines from the library (stderr)
    int device = 0;                // CUDA device ordinal
  };

  struct Summary {    // Optimizer.h:93-99
    double initialCost;
    double finalCost;
    int numTroubledSeqs;
    int largestTroubledSeq;
    int numIterations;
  };

  Optimizer() = default;
  Optimizer(const Optimizer&) = delete;
  Optimizer& operator=(const Optimizer&) = delete;
  ~Optimizer() { releaseBackend(); }

  // ---- addFactor overloads (Optimizer.h:115-175) ------------------------------------------------
  // with precision matrix (row-major dim x dim) and loss
  template <typename Factor, typename SoftLoss, typename... Variables,
            std::enable_if_t<std::is_base_of_v<Loss, SoftLoss>, int> = 0>
  auto addFactor(Factor&& f, const double* precisionMatrix, const SoftLoss& l, Variables&... v) {
    return addImpl(std::forward<Factor>(f), precisionMatrix, &l, v...);
  }
  // with precision matrix, trivial loss
  template <typename Factor, typename Variable0, typename... Variables,
            std::enable_if_t<std::is_base_of_v<VarBase, Variable0>, int> = 0>
  auto addFactor(Factor&& f, const double* precisionMatrix, Variable0& v0, Variables&... v) {
    return addImpl(std::forward<Factor>(f), precisionMatrix, &trivialLoss(), v0, v...);
  }
  // with loss
  template <typename Factor, typename SoftLoss, typename... Variables,
            std::enable_if_t<std::is_base_of_v<Loss, SoftLoss>, int> = 0>
  auto addFactor(Factor&& f, const SoftLoss& l, Variables&... v) {
    return addImpl(std::forward<Factor>(f), nullptr, &l, v...);
  }
  // plain
  template <typename Factor, typename Variable0, typename... Variables,
            std::enable_if_t<std::is_base_of_v<VarBase, Variable0>, int> = 0>
  auto addFactor(Factor&& f, Variable0& v0, Variables&... v) {
    return addImpl(std::forward<Factor>(f), nullptr, &trivialLoss(), v0, v...);
  }

  // ---- registration (Optimizer.h:193-211) -------------------------------------------------------
  template <typename Variable>
  void registerVariable(Variable& v) {
    if (v.index == kUnsetIndex) {
      v.index = (int64_t)paramSizes.size();
      paramSizes.push_back(v.getTangentDim());
      registered_.push_back(&v);
      dirty_ = true;
    }
  }

  void unregisterAllVariables() {
    for (VarBase* v : registered_) v->index = kUnsetIndex;
    registered_.clear();
    paramSizes.clear();
    elimRanges.clear();
    dirty_ = true;
  }

  // every variable referenced by a factor gets an index unless constant: per store, per argument
  // position, per factor in insertion order (Optimizer.cpp:25-39, Factor.h:737-758)
  void registerAllVariables() {
    for (auto& st : factorStores) {
      for (int a = 0; a < st->numVars(); a++) {
        for (int64_t k = 0; k < st->numCosts(); k++) {
          VarBase* v = st->costVar(k, a);
          if (v->index == kUnsetIndex) {
            v->index = (int64_t)paramSizes.size();
            paramSizes.push_back(v->getTangentDim());
            registered_.push_back(v);
            dirty_ = true;
          }
        }
      }
    }
  }

  // make what is registered so far one Schur-eliminated block range (Optimizer.cpp:41-47)
  void registeredVariablesToEliminationRange() {
    if (elimRanges.empty()) elimRanges.push_back(0);
    elimRanges.push_back((int64_t)paramSizes.size());
    dirty_ = true;
  }

  int64_t totalVarDimensionality() const {
    int64_t t = 0;
    for (int64_t s : paramSizes) t += s;
    return t;
  }

  int64_t totalErrorDimensionality() const {
    static const std::map<int32_t, int> err = {{1, 2}, {2, 2}, {4, 9}, {5, 9}, {6, 9}, {7, 23}, {8, 17}, {9, 6},
                                               {11, 23}, {12, 17}, {13, 6}, {15, 3}, {16, 3}};
    int64_t t = 0;
    for (auto& st : factorStores) t += st->numCosts() * err.at(st->abiType());
    return t;
  }

  int64_t variableBackupSize() const {
    int64_t t = 0;
    for (VarBase* v : registered_) t += v->getDataDim();
    return t;
  }

  // ---- compute entry points ---------------------------------------------------------------------
  // Optimizer.h:232-235 / Optimizer.cpp:88-97 -> vb_eval_cost
  double computeCost(bool makeComparableWithStored = false, CostStats* stats = nullptr) {
    ensureBackend(ctx_ ? device_ : 0);   // stay on the device optimize() was last asked to use
    double cost = 0.0;
    vb_cost_stats cs{};
    check(vb_eval_cost(problem_, makeComparableWithStored ? 1 : 0, &cost, &cs));
    if (stats) {
      stats->numTotal = cs.num_total;
      stats->numInvalid = cs.num_invalid;
      stats->numPrevInvalid = cs.num_prev_invalid;
    }
    return cost;
  }

  // the device problem behind this optimizer, built / refreshed from the current graph (for BaSpaCho::Solver of
  // include/baspacho/baspacho/Solver.h: the reference's `solver` member, Optimizer.h:267)
  vb_problem* backendProblem(int device = 0) {
    ensureBackend(ctx_ ? device_ : device);
    return problem_;
  }

  // ---- covariances and marginal problems (Optimizer.h:101-106,294-326; Optimizer.cpp:422-696) -----------------------
  // Eigen::MatrixXd of the reference -> a plain row-major matrix
  struct DenseMatrix {
    int64_t rows = 0, cols = 0;
    std::vector<double> data;
    double& operator()(int64_t r, int64_t c) { return data[(size_t)(r * cols + c)]; }
    double operator()(int64_t r, int64_t c) const { return data[(size_t)(r * cols + c)]; }
  };
  struct MarginalProblem {   // Optimizer.h:101-106
    DenseMatrix H;
    std::vector<double> b;
    double cost = 0.0;
    std::vector<int64_t> offsets;
  };

  // returns a list of pairs {matrix, offsets}: joint covariance and offsets for each list of registration indices
  // (Optimizer.cpp:503-612 -> vb_compute_joint_covariances: one linearisation + factorisation for all lists)
  std::vector<std::pair<DenseMatrix, std::vector<int64_t>>> computeJointCovariances(const Settings& s,
                                                                                     const std::vector<std::vector<int64_t>>& indicesLists) {
    std::vector<std::pair<DenseMatrix, std::vector<int64_t>>> out;
    if (indicesLists.empty()) return out;
    ensureBackend(s.device);
    std::vector<int64_t> ptr{0}, flat;
    int64_t total = 0;
    for (const auto& list : indicesLists) {
      std::vector<int64_t> off{0};
      for (int64_t i : list) {
        if (i < 0 || i >= (int64_t)registered_.size()) throw std::runtime_error("small_thing: taking index from UNREGISTERED variable?");
        off.push_back(off.back() + registered_[(size_t)i]->getTangentDim());
        flat.push_back(i);
      }
      ptr.push_back((int64_t)flat.size());
      DenseMatrix m;
      m.rows = m.cols = off.back();
      total += m.rows * m.cols;
      out.emplace_back(std::move(m), std::move(off));
    }
    std::vector<double> cov((size_t)total);
    std::vector<int64_t> covPtr(indicesLists.size() + 1);
    check(vb_compute_joint_covariances(problem_, s.damping, (int64_t)indicesLists.size(), ptr.data(), flat.data(), cov.data(), total,
                                       covPtr.data(), nullptr));
    for (size_t l = 0; l < out.size(); l++) out[l].first.data.assign(cov.begin() + covPtr[l], cov.begin() + covPtr[l + 1]);
    return out;
  }
  std::vector<std::pair<DenseMatrix, std::vector<int64_t>>> computeJointCovariances(const std::vector<std::vector<int64_t>>& indicesLists) {
    return computeJointCovariances(Settings{}, indicesLists);
  }
  // compute covariance block for variables in list `indices`, one per variable (Optimizer.cpp:613-696)
  std::vector<DenseMatrix> computeCovariances(const Settings& s, const std::vector<int64_t>& indices) {
    std::vector<std::vector<int64_t>> lists;
    for (int64_t i : indices) lists.push_back({i});
    std::vector<DenseMatrix> out;
    for (auto& pr : computeJointCovariances(s, lists)) out.push_back(std::move(pr.first));
    return out;
  }
  std::vector<DenseMatrix> computeCovariances(const std::vector<int64_t>& indices) { return computeCovariances(Settings{}, indices); }
  // compute big marginal block H and error b (Optimizer.cpp:422-494 -> vb_compute_marginal_problem)
  MarginalProblem computeMarginalProblem(const Settings& s, const std::vector<int64_t>& indices) {
    ensureBackend(s.device);
    MarginalProblem m;
    m.offsets.assign(indices.size() + 1, 0);
    for (size_t k = 0; k < indices.size(); k++) {
      if (indices[k] < 0 || indices[k] >= (int64_t)registered_.size()) throw std::runtime_error("small_thing: taking index from UNREGISTERED variable?");
      m.offsets[k + 1] = m.offsets[k] + registered_[(size_t)indices[k]]->getTangentDim();
    }
    const int64_t D = m.offsets.back();
    m.H.rows = m.H.cols = D;
    m.H.data.assign((size_t)(D * D), 0.0);
    m.b.assign((size_t)D, 0.0);
    check(vb_compute_marginal_problem(problem_, s.damping, (int64_t)indices.size(), indices.data(), m.H.data.data(), D * D, m.b.data(), D, &m.cost,
                                      nullptr, nullptr));
    return m;
  }
  MarginalProblem computeMarginalProblem(const std::vector<int64_t>& indices) { return computeMarginalProblem(Settings{}, indices); }

  Summary optimize() { return optimize(Settings{}); }

  // Optimizer.cpp:768-1106 -> vb_optimize (same control flow, device resident), then the values
  // are written back into the user's Variables
  Summary optimize(const Settings& s) {
    ensureBackend(s.device);
    vb_settings vs;
    vb_settings_default(&vs);
    vs.damping = s.damping;
    vs.damping_adjust_on_fail = s.dampingAdjustOnFail;
    vs.damping_adjust_on_good = s.dampingAdjustOnGoodStep;
    vs.damping_adjust_on_average = s.dampingAdjustOnAverageStep;
    vs.damping_min = s.dampingMin;
    vs.damping_max = s.dampingMax;
    vs.min_relative_cost_reduction = s.minRelativeCostReduction;
    vs.step_factor_decrease = s.stepFactorDecrease;
    vs.min_step_factor_for_good = s.minStepFactorForGood;
    vs.absolute_cost_tolerance = s.absoluteCostTolerance;
    vs.relative_cost_tolerance = s.relativeCostTolerance;
    vs.variables_tolerance = s.variablesTolerance;
    vs.max_step_factor_attempts = s.maxStepFactorAttempts;
    vs.try_sub_step = s.trySubStep ? 1 : 0;
    vs.stop_if_no_improvement_for = s.stopIfNoImprovementFor;
    vs.distance_from_troubled_iteration = s.distanceFromTroubledIteration;
    vs.max_num_iterations = s.maxNumIterations;
    vs.verbose = s.verbose ? 1 : 0;
    vs.solver_type = (int32_t)s.solverType;   // same numbering as VB_SOLVER_*; the library rejects anything else
    vs.pcg_max_iterations = s.pcgMaxIterations;
    vs.pcg_desired_residual = s.pcgDesiredResidual;
    vb_summary out{};
    struct Hook { Optimizer* self; const Settings* s; std::string err; } hk{this, &s, {}};
    vb_optimize_hooks hooks{};
    if (s.preStepCallback) {
      hooks.pre_step_user = &hk;
      hooks.sync_kinds = (1u << VB_VAR_IMUCALIB) | (1u << VB_VAR_CAM) | (1u << VB_VAR_S2);
      hooks.host_vars = graph_->data.data();
      hooks.pre_step = [](void* u, int32_t it) {
        Hook& h = *static_cast<Hook*>(u);
        if (!h.err.empty()) return;
        try {
          ExportedGraph& g = *h.self->graph_;
          for (size_t i = 0; i < g.pool.size(); i++)   // the kinds the library just synchronised into host_vars
            if (g.index[i] >= 0 && (g.kind[i] == VB_VAR_IMUCALIB || g.kind[i] == VB_VAR_CAM || g.kind[i] == VB_VAR_S2))
              g.pool[i]->setData(g.data.data() + g.off[i]);
          h.s->preStepCallback((int)it);
          h.self->pushStoreData();
        } catch (const std::exception& ex) { h.err = ex.what(); }   // no exception may cross the C ABI
      };
    }
    if (s.log) {
      hooks.log_user = &hk;
      hooks.log = [](void* u, const char* text) { static_cast<Hook*>(u)->s->log(text); };
    }
    check(vb_optimize_ex(problem_, &vs, &hooks, &out));
    if (!hk.err.empty()) throw std::runtime_error("small_thing: preStepCallback threw: " + hk.err);
    pullValues();
    return Summary{out.initial_cost, out.final_cost, out.num_troubled_seqs, out.largest_troubled_seq, out.num_iterations};
  }

  // ---- the graph as handed to the C ABI (also what tests inspect) -------------------------------
  // Pool ids: registered variables in registration order, then constants in order of first use.
  std::unique_ptr<ExportedGraph> exportGraph() {
    registerAllVariables();
    auto g = std::make_unique<ExportedGraph>();
    std::unordered_map<const VarBase*, int32_t> id;
    auto varId = [&](VarBase* v) -> int32_t {
      auto it = id.find(v);
      if (it != id.end()) return it->second;
      const int32_t k = (int32_t)g->pool.size();
      id.emplace(v, k);
      g->pool.push_back(v);
      g->kind.push_back(v->kind());
      g->tdim.push_back(v->getTangentDim());
      g->flags.push_back(v->flags());
      g->index.push_back(v->index);
      g->aux.push_back(v->aux());
      g->off.push_back((int64_t)g->data.size());
      g->data.resize(g->data.size() + (size_t)v->getDataDim());
      v->getData(g->data.data() + g->off.back());
      return k;
    };
    for (VarBase* v : registered_) varId(v);
    int64_t numPoints = 0;
    if (!elimRanges.empty()) {
      if (elimRanges.size() != 2 || elimRanges[0] != 0)
        throw std::runtime_error("small_thing: the device path supports one elimination range starting at 0");
      numPoints = elimRanges[1];
      for (int64_t i = 0; i < numPoints; i++)
        if (registered_[(size_t)i]->kind() != VB_VAR_VEC3)
          throw std::runtime_error("small_thing: the elimination range must hold 3-D point variables only");
    }
    const size_t ns = factorStores.size();
    g->storeVars.resize(ns);
    g->storeConsts.resize(ns);
    g->storeAux.resize(ns);
    g->stores.resize(ns);
    for (size_t s = 0; s < ns; s++) {
      FactorStoreBase& st = *factorStores[s];
      const int nv = st.numVars();
      const int64_t n = st.numCosts();
      auto& sv = g->storeVars[s];
      sv.resize((size_t)(n * nv));
      for (int64_t k = 0; k < n; k++)
        for (int a = 0; a < nv; a++) sv[(size_t)(k * nv + a)] = varId(st.costVar(k, a));
      g->storeConsts[s].resize((size_t)(n * st.constsPerFactor()));
      st.fillConsts(g->storeConsts[s].data());
      st.fillAux(g->storeAux[s]);
      vb_factor_store& o = g->stores[s];
      o.type = st.abiType();
      o.nvars = nv;
      o.count = n;
      o.vars = sv.data();
      o.consts_per_factor = st.constsPerFactor();
      o.consts = g->storeConsts[s].data();
      o.loss_a = st.loss ? st.loss->abiA() : 0.0;
      o.loss_k = st.loss ? st.loss->abiK() : 0.0;
      o.aux_len = (int64_t)g->storeAux[s].size();
      o.aux = g->storeAux[s].empty() ? nullptr : g->storeAux[s].data();
    }
    vb_graph_desc& d = g->desc;
    d.num_vars = (int64_t)g->pool.size();
    d.var_kind = g->kind.data();
    d.var_tdim = g->tdim.data();
    d.var_flags = g->flags.data();
    d.var_index = g->index.data();
    d.var_off = g->off.data();
    d.var_aux = g->aux.data();
    d.var_data_len = (int64_t)g->data.size();
    d.var_data = g->data.data();
    d.num_points = numPoints;
    d.num_stores = (int32_t)ns;
    d.stores = g->stores.data();
    return g;
  }

  // introspection, as the reference exposes its stores (Histograms.cpp, Stats.cpp walk them)
  std::vector<std::unique_ptr<FactorStoreBase>> factorStores;
  std::vector<int64_t> paramSizes;
  std::vector<int64_t> elimRanges;
  const std::vector<VarBase*>& registeredVariables() const { return registered_; }

 private:
  static const TrivialLoss& trivialLoss() {
    static const TrivialLoss l;
    return l;
  }

  static void check(int status) {
    if (status != VB_OK) throw std::runtime_error(std::string("viba_b200: ") + vb_last_error());
  }

  template <typename Factor, typename... Variables>
  auto addImpl(Factor&& f, const double* precision, const Loss* loss, Variables&... v) {
    using F = std::decay_t<Factor>;
    static_assert((std::is_base_of_v<VarBase, Variables> && ...), "addFactor: arguments after the loss must be Variables");
    constexpr int nv = (int)sizeof...(Variables);
    const int32_t type = F::abiType(nv);
    if (type < 0) throw std::runtime_error(std::string("addFactor: ") + F::name() + " does not take " + std::to_string(nv) + " variables");
    if (F::kHasPrecision != (precision != nullptr))
      throw std::runtime_error(std::string("addFactor: ") + F::name() + (F::kHasPrecision ? " needs" : " does not take") + " a precision matrix");
    VarBase* args[] = {static_cast<VarBase*>(&v)...};
    const int32_t* kinds = F::kinds(nv);
    for (int a = 0; a < nv; a++)
      if (args[a]->kind() != kinds[a])
        throw std::runtime_error(std::string("addFactor: ") + F::name() + ": wrong variable type at argument " + std::to_string(a));
    // one store per (payload type, arity, loss object type), as FactorStore<F, HasPrecision, SoftLoss, Variables...>
    const StoreKey key{std::type_index(typeid(F)), nv, std::type_index(typeid(*loss))};
    auto it = storeOf_.find(key);
    if (it == storeOf_.end()) {
      it = storeOf_.emplace(key, factorStores.size()).first;
      factorStores.emplace_back(std::make_unique<FactorStore<F>>(nv));
      factorStores.back()->loss = loss;
    }
    auto* store = static_cast<FactorStore<F>*>(factorStores[it->second].get());
    if (store->loss != loss && (store->loss->abiA() != loss->abiA() || store->loss->abiK() != loss->abiK()))
      throw std::runtime_error(std::string("addFactor: ") + F::name() + ": one store, two different losses");
    store->factors.push_back(std::forward<Factor>(f));
    for (VarBase* a : args) store->vars.push_back(a);
    if constexpr (F::kHasPrecision) {
      store->precisionDim = F::kPrecisionDim;
      store->precision.insert(store->precision.end(), precision, precision + F::kPrecisionDim * F::kPrecisionDim);
    }
    dirty_ = true;
    return std::pair<FactorStore<F>*, int64_t>(store, (int64_t)store->factors.size() - 1);
  }

  // (re)build the device problem when the graph changed, else refresh values and per-store data
  void ensureBackend(int device) {
    if (ctx_ && device_ != device) releaseBackend();
    if (!ctx_) {
      check(vb_ctx_create(device, &ctx_));
      device_ = device;
      dirty_ = true;
    }
    graph_ = exportGraph();   // current host values, constants, losses, user records
    if (dirty_ || !problem_) {
      if (problem_) {
        vb_problem_destroy(problem_);
        problem_ = nullptr;
      }
      check(vb_problem_create(ctx_, &graph_->desc, &problem_));
      dirty_ = false;
    } else {
      check(vb_upload_vars(problem_, graph_->data.data(), (int64_t)graph_->data.size()));
      for (size_t s = 0; s < graph_->stores.size(); s++) {
        if (!graph_->storeConsts[s].empty())
          check(vb_upload_store_consts(problem_, (int32_t)s, graph_->storeConsts[s].data(), (int64_t)graph_->storeConsts[s].size()));
        if (!graph_->storeAux[s].empty())
          check(vb_upload_store_aux(problem_, (int32_t)s, graph_->storeAux[s].data(), (int64_t)graph_->storeAux[s].size()));
      }
    }
  }

  // user-owned per-store data (factor constants, preintegration records, rolling-shutter tables) -> device
  void pushStoreData() {
    for (size_t s = 0; s < factorStores.size(); s++) {
      FactorStoreBase& st = *factorStores[s];
      if (!graph_->storeConsts[s].empty()) {
        st.fillConsts(graph_->storeConsts[s].data());
        check(vb_upload_store_consts(problem_, (int32_t)s, graph_->storeConsts[s].data(), (int64_t)graph_->storeConsts[s].size()));
      }
      const size_t before = graph_->storeAux[s].size();
      st.fillAux(graph_->storeAux[s]);
      if (graph_->storeAux[s].size() != before) throw std::runtime_error("small_thing: a callback changed the size of a store's records");
      if (before) check(vb_upload_store_aux(problem_, (int32_t)s, graph_->storeAux[s].data(), (int64_t)before));
    }
  }

  void pullValues() {
    check(vb_download_vars(problem_, graph_->data.data(), (int64_t)graph_->data.size()));
    for (size_t i = 0; i < graph_->pool.size(); i++)
      if (graph_->index[i] >= 0) graph_->pool[i]->setData(graph_->data.data() + graph_->off[i]);
  }

  void releaseBackend() {
    if (problem_) vb_problem_destroy(problem_);
    if (ctx_) vb_ctx_destroy(ctx_);
    problem_ = nullptr;
    ctx_ = nullptr;
  }

  struct StoreKey {
    std::type_index f;
    int nvars;
    std::type_index loss;
    bool operator<(const StoreKey& o) const {
      if (f != o.f) return f < o.f;
      if (nvars != o.nvars) return nvars < o.nvars;
      return loss < o.loss;
    }
  };
  std::map<StoreKey, size_t> storeOf_;
  std::vector<VarBase*> registered_;   // variableStores of the reference, in index order
  std::unique_ptr<ExportedGraph> graph_;
  vb_ctx* ctx_ = nullptr;
  vb_problem* problem_ = nullptr;
  int device_ = 0;
  bool dirty_ = true;
};

}  // namespace small_thing
