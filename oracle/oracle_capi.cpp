// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
// C entry points over the oracle so that tests/ (ctypes) and bench.py's cpu_baseline /
// --impl reference legs can drive it.  Never linked into the product library.
#include <dlfcn.h>
#include <cstring>
#include <string>
#include <chrono>
#include "lm.hpp"
#include "rs_refresh.hpp"
#include "preint.hpp"
#include "refine.hpp"

using namespace orc;

struct RsSource {   // raw IMU-0 samples + per-rig windows of one rolling-shutter store
  int store = -1;
  ImuSamples meas;
  std::vector<int64_t> midUs;
  std::vector<int32_t> calibVar;
  int64_t halfUs = 0;
  int32_t gravityVar = -1;   // re-read at every refresh (InitCalibration.cpp:317), -1: keep the stored value
};

struct PreintSource {   // raw samples of one IMU + the integration interval of every factor of one inertial store
  int store = -1;
  ImuSamples meas;
  std::vector<int64_t> startUs, endUs;
  std::vector<int32_t> calibVar;
  double gyroVar[3], accelVar[3];
  std::vector<double> cov, omegaAtEnd;   // per factor: 81 / 3 doubles, filled by the refresh
};

struct OrcHandle {
  std::vector<RsSource> rs;
  std::vector<PreintSource> preint;
  Engine eng;
  SchurSkylineSolver sol;
  bool analyzed = false;
  std::vector<double> hessF;
  std::string err;
};

static thread_local std::string g_err;

#define ORC_TRY try {
#define ORC_CATCH                                   \
  }                                                 \
  catch (const std::exception& ex) {                \
    g_err = ex.what();                              \
    return 3;                                       \
  }                                                 \
  return 0;

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }

int orc_create(const vb_graph_desc* g, OrcHandle** out) {
  ORC_TRY
  auto* h = new OrcHandle();
  h->eng.P.load(g);
  h->eng.symbolic();
  *out = h;
  ORC_CATCH
}
int orc_destroy(OrcHandle* h) { delete h; return 0; }

int64_t orc_order(OrcHandle* h) { return h->eng.order; }
int64_t orc_data_size(OrcHandle* h) { return h->eng.dataSize; }
int64_t orc_num_blocks(OrcHandle* h) { return h->eng.P.nblocks; }
int64_t orc_nnz_blocks(OrcHandle* h) { return (int64_t)h->eng.colInd.size(); }
// block CSR of Optimizer::initSolver (Optimizer.cpp:174-197): rowPtr[n+1], colInd[nnz], blkOff[nnz]
int orc_get_structure(OrcHandle* h, int64_t* rowPtr, int64_t* colInd, int64_t* blkOff, int64_t* paramStart) {
  const Engine& e = h->eng;
  if (rowPtr) std::memcpy(rowPtr, e.rowPtr.data(), e.rowPtr.size() * 8);
  if (colInd) std::memcpy(colInd, e.colInd.data(), e.colInd.size() * 8);
  if (blkOff) std::memcpy(blkOff, e.blkOff.data(), e.blkOff.size() * 8);
  if (paramStart) std::memcpy(paramStart, e.paramStart.data(), e.paramStart.size() * 8);
  return 0;
}

int orc_eval_grad_hess(OrcHandle* h, int with_hess, int update_cache, int dont_retry, int which, double* cost) {
  ORC_TRY
  Engine& e = h->eng;
  std::vector<double>& g = which == 0 ? e.grad : e.gradNew;
  std::fill(g.begin(), g.end(), 0.0);
  if (with_hess) std::fill(e.hess.begin(), e.hess.end(), 0.0);
  const double c = e.computeGradHess(g.data(), with_hess ? e.hess.data() : nullptr, update_cache != 0, dont_retry != 0);
  if (cost) *cost = c;
  ORC_CATCH
}
int orc_eval_cost(OrcHandle* h, int comparable, double* cost, vb_cost_stats* stats) {
  ORC_TRY
  CostStats cs;
  const double c = h->eng.computeCost(comparable != 0, &cs);
  if (cost) *cost = c;
  if (stats) { stats->num_total = cs.numTotal; stats->num_invalid = cs.numInvalid; stats->num_prev_invalid = cs.numPrevInvalid; }
  ORC_CATCH
}
int orc_damp(OrcHandle* h, double lambda) { h->eng.addDamping(lambda); return 0; }

// factor (on a private copy of hess) then step = H^-1 grad
int orc_factor(OrcHandle* h) {
  ORC_TRY
  if (!h->analyzed) { h->sol.analyze(h->eng); h->analyzed = true; }
  h->hessF = h->eng.hess;
  h->sol.factor(h->hessF.data());
  ORC_CATCH
}
int orc_solve(OrcHandle* h, int which) {
  ORC_TRY
  Engine& e = h->eng;
  if (which == 0) h->sol.solveOnly(h->hessF.data(), e.grad.data(), e.step.data());
  else { std::vector<double> g2 = e.gradNew; h->sol.solveOnly(h->hessF.data(), g2.data(), e.gradNew.data()); }
  ORC_CATCH
}
int orc_scale_vec(OrcHandle* h, int which, double s) {
  for (double& x : (which == 0 ? h->eng.step : h->eng.gradNew)) x *= s;
  return 0;
}
int orc_apply_step(OrcHandle* h, int which, double* ratios) {
  double r[3];
  h->eng.applyStep(which == 0 ? h->eng.step.data() : h->eng.gradNew.data(), r);
  if (ratios) { ratios[0] = r[0]; ratios[1] = r[1]; ratios[2] = r[2]; }
  return 0;
}
int orc_backup(OrcHandle* h) { h->eng.backup(); return 0; }
int orc_restore(OrcHandle* h) { h->eng.restore(); return 0; }

// which: 0 grad, 1 step, 2 gradNew, 3 hess, 4 var data (registration order for vectors)
int orc_get(OrcHandle* h, int which, double* out, int64_t n) {
  Engine& e = h->eng;
  const std::vector<double>* v = which == 0 ? &e.grad : which == 1 ? &e.step : which == 2 ? &e.gradNew : which == 3 ? &e.hess : &e.P.data;
  if ((int64_t)v->size() != n) { g_err = "orc_get: size mismatch"; return 1; }
  std::memcpy(out, v->data(), n * 8);
  return 0;
}
int orc_set(OrcHandle* h, int which, const double* in, int64_t n) {
  Engine& e = h->eng;
  std::vector<double>* v = which == 0 ? &e.grad : which == 1 ? &e.step : which == 2 ? &e.gradNew : which == 3 ? &e.hess : &e.P.data;
  if ((int64_t)v->size() != n) { g_err = "orc_set: size mismatch"; return 1; }
  std::memcpy(v->data(), in, n * 8);
  return 0;
}
int orc_get_cache(OrcHandle* h, int store, double* out, int64_t n) {
  const auto& c = h->eng.P.stores[store].cache;
  if ((int64_t)c.size() != n) return 1;
  std::memcpy(out, c.data(), n * 8);
  return 0;
}
int orc_set_store_aux(OrcHandle* h, int store, const double* aux, int64_t n) {
  h->eng.P.stores[store].aux.assign(aux, aux + n);
  return 0;
}

// per-factor residual + Jacobians: r[E], J = concatenated row-major E x d_i for every variable
// (constants included, so numeric checks can cover them); returns validity in *valid
int orc_eval_factor(OrcHandle* h, int store, int64_t k, double* r, double* J, int* valid) {
  ORC_TRY
  const Engine& e = h->eng;
  const Store& st = e.P.stores[store];
  const int E = factorErrorSize(st.type);
  double* Jp[kMaxVars];
  int64_t o = 0;
  for (int i = 0; i < st.nvars; i++) {
    Jp[i] = J ? J + o : nullptr;
    o += (int64_t)E * e.P.tdim[st.vars[k * st.nvars + i]];
  }
  *valid = e.evalFactor(st, k, r, J ? Jp : nullptr) ? 1 : 0;
  ORC_CATCH
}

int orc_optimize(OrcHandle* h, const vb_settings* s, vb_summary* out, double* tracePrev, double* traceNew, int traceCap) {
  ORC_TRY
  if (!h->analyzed) { h->sol.analyze(h->eng); h->analyzed = true; }
  LmTrace tr;
  optimize(h->eng, h->sol, *s, out, &tr);
  for (int i = 0; i < traceCap && i < (int)tr.prevCost.size(); i++) {
    if (tracePrev) tracePrev[i] = tr.prevCost[i];
    if (traceNew) traceNew[i] = tr.newCost[i];
  }
  ORC_CATCH
}
// same run with the full per-iteration trace (the fields of the reference's log line, Optimizer.cpp:1058-1076,
// and the branches taken) and an optional Settings::preStepCallback stand-in: `cb(user, iteration)` may
// rewrite store aux data / variables through orc_set_store_aux / orc_set before the linearisation
typedef void (*orc_pre_step_fn)(void* user, int32_t iteration);
int orc_optimize_ex(OrcHandle* h, const vb_settings* s, vb_summary* out, vb_iter_trace* trace, int traceCap, int* traceLen,
                    orc_pre_step_fn cb, void* user) {
  ORC_TRY
  if (!h->analyzed) { h->sol.analyze(h->eng); h->analyzed = true; }
  LmTrace tr;
  std::function<void(int)> pre;
  if (cb) pre = [&](int it) { cb(user, it); };
  optimize(h->eng, h->sol, *s, out, &tr, pre);
  int n = 0;
  for (; n < traceCap && n < (int)tr.full.size(); n++) if (trace) trace[n] = tr.full[n];
  if (traceLen) *traceLen = n;
  ORC_CATCH
}
static void preintRefresh(OrcHandle* h, int store) {
  for (PreintSource& src : h->preint) {
    if (store >= 0 && src.store != store) continue;
    Problem& P = h->eng.P;
    double* recs = P.stores[src.store].aux.data();
    const int threads = h->eng.threads;
    std::string failure;
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads) if (threads > 1)
    for (int64_t k = 0; k < (int64_t)src.startUs.size(); k++) {
      if (src.calibVar[k] < 0) continue;
      try {
        const int32_t v = src.calibVar[k];
        const PreintResult r = preintCompute(src.meas, P.data.data() + P.off[v], P.flags[v], src.gyroVar, src.accelVar, src.startUs[k], src.endUs[k]);
        std::memcpy(recs + k * kPreintRec, r.rec, sizeof r.rec);
        std::memcpy(src.cov.data() + 81 * k, r.cov, sizeof r.cov);
        std::memcpy(src.omegaAtEnd.data() + 3 * k, r.omegaAtEnd, sizeof r.omegaAtEnd);
      } catch (const std::exception& ex) {
#pragma omp critical
        failure = ex.what();
      }
    }
    if (!failure.empty()) throw std::runtime_error(failure);
  }
}
static void rsRefresh(OrcHandle* h, int store) {
  for (const RsSource& src : h->rs) {
    if (store >= 0 && src.store != store) continue;
    Problem& P = h->eng.P;
    if (src.gravityVar >= 0) for (int i = 0; i < 3; i++) P.stores[src.store].aux[i] = P.data[P.off[src.gravityVar] + i];
    double* tables = P.stores[src.store].aux.data() + 4;
    const int threads = h->eng.threads;
    std::string failure;   // exceptions must not leave the OpenMP region
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads) if (threads > 1)
    for (int64_t r = 0; r < (int64_t)src.midUs.size(); r++) {
      if (src.calibVar[r] < 0) continue;
      try {
        rsCompute(src.meas, P.data.data() + P.off[src.calibVar[r]], src.midUs[r], src.halfUs, tables + r * kRsStride);
      } catch (const std::exception& ex) {
#pragma omp critical
        failure = ex.what();
      }
    }
    if (!failure.empty()) throw std::runtime_error(failure);
  }
}

// SingleSessionAdapter::updateRollingShutterData (InitCalibration.cpp:316-325): same arguments as vb_rs_set_imu
int orc_rs_set_imu(OrcHandle* h, int store, int64_t n_rigs, const int64_t* rig_mid_time_us, const int32_t* rig_calib_var,
                   int64_t half_length_us, int64_t n_samples, const int64_t* sample_time_ns, const float* gyro, const float* accel) {
  ORC_TRY
  RsSource src;
  src.store = store;
  src.halfUs = half_length_us;
  src.midUs.assign(rig_mid_time_us, rig_mid_time_us + n_rigs);
  src.calibVar.assign(rig_calib_var, rig_calib_var + n_rigs);
  src.meas.timeNs.assign(sample_time_ns, sample_time_ns + n_samples);
  src.meas.gyro.assign(gyro, gyro + 3 * n_samples);
  src.meas.accel.assign(accel, accel + 3 * n_samples);
  if ((int64_t)h->eng.P.stores.at(store).aux.size() != 4 + n_rigs * kRsStride) throw std::runtime_error("rig count does not match the store's tables");
  h->rs.push_back(std::move(src));
  h->eng.preStep = [h]() { preintRefresh(h, -1); rsRefresh(h, -1); };
  ORC_CATCH
}
// Settings::solverType for orc_factor / orc_solve (orc_optimize takes it from its settings); last PCG run's statistics
int orc_set_linear_solver(OrcHandle* h, int solver_type, int pcg_max_iterations, double pcg_desired_residual) {
  ORC_TRY
  h->sol.setLinearSolver(solver_type, pcg_max_iterations, pcg_desired_residual);
  ORC_CATCH
}
int orc_pcg_last(OrcHandle* h, int* iterations, double* residual) {
  *iterations = h->sol.pcgLastIterations; *residual = h->sol.pcgLastResidual;
  return 0;
}
// PCG on a dense symmetric system, identity (0) or block-Jacobi (1) preconditioner: the reduced system of the reference's
// own PCG test (lib/small_thing/tests/TestPCG.cpp:29-131), whose recipe tests/test_pcg.py restates
int orc_pcg_dense(int64_t n, const double* A, const double* b, const int32_t* block_size, int64_t n_blocks, int precond,
                  double desired_residual, int max_iterations, double* x, int* iterations, double* residual) {
  try {
    const orc::PcgResult r = orc::pcgDense(n, A, b, block_size, n_blocks, precond, desired_residual, max_iterations, x);
    *iterations = r.numIterations; *residual = r.relativeResidual;
    return 0;
  } catch (const std::exception&) { return 1; }
}
// refinePoints (viba/problem/PointRefinement.cpp:171-196): stats6 = {totalStartCost, totalEndCost, nPoints, nFailures,
// nTotalIterations, nAtLeastOneIt}; per_point (may be null) = {startCost, endCost, nIts} per refined landmark
int orc_refine_points(OrcHandle* h, double* stats6, double* per_point) {
  ORC_TRY
  const RefineStats r = refinePoints(h->eng, per_point);
  stats6[0] = r.totalStartCost; stats6[1] = r.totalEndCost; stats6[2] = (double)r.nPoints; stats6[3] = (double)r.nFailures;
  stats6[4] = (double)r.nTotalIterations; stats6[5] = (double)r.nAtLeastOneIt;
  ORC_CATCH
}
// refineTriangulationResult (viba/single_session/Triangulation.cpp:99-161) over every landmark: per_point = {numInliers,
// last cost, status} per refined landmark; inlier_flags (may be null) = one byte per factor, the visual stores
// concatenated in graph order; *failures = landmarks whose 3x3 system was not positive definite
int orc_refine_triangulation(OrcHandle* h, double outlier_threshold, int skip_outliers, int max_num_iterations, double huber_radius,
                             double* per_point, uint8_t* inlier_flags, int64_t* failures) {
  ORC_TRY
  std::vector<std::vector<uint8_t>> fl;
  const int64_t f = refineTriangulation(h->eng, outlier_threshold, skip_outliers != 0, max_num_iterations, huber_radius, per_point, &fl);
  if (failures) *failures = f;
  if (inlier_flags) { int64_t o = 0; for (const auto& v : fl) { std::copy(v.begin(), v.end(), inlier_flags + o); o += (int64_t)v.size(); } }
  ORC_CATCH
}
// SingleSessionAdapter::regenerateAllPreintegrationsFromImuMeasurements (InertialFactors.cpp:45-70): same arguments as vb_preint_set_imu
int orc_preint_set_imu(OrcHandle* h, int store, int64_t n_factors, const int64_t* start_us, const int64_t* end_us, const int32_t* calib_var,
                       int64_t n_samples, const int64_t* sample_time_ns, const float* gyro, const float* accel, const double* gyro_var3,
                       const double* accel_var3) {
  ORC_TRY
  PreintSource src;
  src.store = store;
  src.startUs.assign(start_us, start_us + n_factors);
  src.endUs.assign(end_us, end_us + n_factors);
  src.calibVar.assign(calib_var, calib_var + n_factors);
  src.meas.timeNs.assign(sample_time_ns, sample_time_ns + n_samples);
  src.meas.gyro.assign(gyro, gyro + 3 * n_samples);
  src.meas.accel.assign(accel, accel + 3 * n_samples);
  for (int i = 0; i < 3; i++) { src.gyroVar[i] = gyro_var3[i]; src.accelVar[i] = accel_var3[i]; }
  src.cov.assign(81 * n_factors, 0.0);
  src.omegaAtEnd.assign(3 * n_factors, 0.0);
  if ((int64_t)h->eng.P.stores.at(store).aux.size() != n_factors * kPreintRec) throw std::runtime_error("factor count does not match the store's records");
  h->preint.push_back(std::move(src));
  h->eng.preStep = [h]() { preintRefresh(h, -1); rsRefresh(h, -1); };
  ORC_CATCH
}
int orc_preint_refresh(OrcHandle* h, int store) {
  ORC_TRY
  preintRefresh(h, store);
  ORC_CATCH
}
int orc_preint_get_cov(OrcHandle* h, int store, double* cov, double* omega_at_end, int64_t n_factors) {
  ORC_TRY
  for (const PreintSource& src : h->preint) {
    if (src.store != store) continue;
    if ((int64_t)src.startUs.size() != n_factors) throw std::runtime_error("factor count mismatch");
    if (cov) std::memcpy(cov, src.cov.data(), src.cov.size() * 8);
    if (omega_at_end) std::memcpy(omega_at_end, src.omegaAtEnd.data(), src.omegaAtEnd.size() * 8);
    return 0;
  }
  throw std::runtime_error("store has no IMU source");
  ORC_CATCH
}
// computePreIntegration (PreIntegration.cpp:136-275) on raw samples, standalone
int orc_preint_compute(int64_t n_samples, const int64_t* sample_time_ns, const float* gyro, const float* accel, const double* calib32,
                       uint32_t flags, const double* gyro_var3, const double* accel_var3, int64_t start_us, int64_t end_us, double* rec250,
                       double* cov81, double* omega3, int32_t* steps) {
  ORC_TRY
  ImuSamples m;
  m.timeNs.assign(sample_time_ns, sample_time_ns + n_samples);
  m.gyro.assign(gyro, gyro + 3 * n_samples);
  m.accel.assign(accel, accel + 3 * n_samples);
  const PreintResult r = preintCompute(m, calib32, flags, gyro_var3, accel_var3, start_us, end_us);
  std::memcpy(rec250, r.rec, sizeof r.rec);
  if (cov81) std::memcpy(cov81, r.cov, sizeof r.cov);
  if (omega3) std::memcpy(omega3, r.omegaAtEnd, sizeof r.omegaAtEnd);
  if (steps) *steps = r.steps;
  ORC_CATCH
}
// pieces, for the reference's unit-test recipes: integrate + Jacobian (MotionIntegral.cpp:162-226), combineJacs (:52-75),
// getCompensatedImuMeasurementAndJac (CompensateJac.cpp:146-249), combine / differentiate / integrate(interp)
int orc_integrate_jac(const double* gyro3, const double* accel3, double dt, double* out11, double* J54) {
  ORC_TRY
  double J[54];
  const Rvp r = integrateJac(gyro3, accel3, dt, J);
  const Rvp r0 = integrate(gyro3, accel3, dt);
  for (int i = 0; i < 4; i++) if (r.q[i] != r0.q[i]) throw std::runtime_error("integrate with and without Jacobian differ");
  for (int i = 0; i < 4; i++) out11[i] = r.q[i];
  for (int i = 0; i < 3; i++) { out11[4 + i] = r.dV[i]; out11[7 + i] = r.dP[i]; }
  out11[10] = r.dt;
  if (J54) std::memcpy(J54, J, sizeof J);
  ORC_CATCH
}
static Rvp rvpFrom(const double* a) { Rvp r; for (int i = 0; i < 4; i++) r.q[i] = a[i]; for (int i = 0; i < 3; i++) { r.dV[i] = a[4 + i]; r.dP[i] = a[7 + i]; } r.dt = a[10]; return r; }
static void rvpTo(const Rvp& r, double* a) { for (int i = 0; i < 4; i++) a[i] = r.q[i]; for (int i = 0; i < 3; i++) { a[4 + i] = r.dV[i]; a[7 + i] = r.dP[i]; } a[10] = r.dt; }
int orc_combine(const double* a11, const double* b11, double* c11) {
  ORC_TRY
  rvpTo(combine(rvpFrom(a11), rvpFrom(b11)), c11);
  ORC_CATCH
}
int orc_uncombine_left(const double* c11, const double* a11, double* b11) {
  ORC_TRY
  rvpTo(uncombineLeft(rvpFrom(c11), rvpFrom(a11)), b11);
  ORC_CATCH
}
int orc_combine_jacs(const double* a11, const double* b11, const double* aJ, const double* bJ, int n, double* c11, double* cJ) {
  ORC_TRY
  rvpTo(combineJacs(rvpFrom(a11), rvpFrom(b11), aJ, bJ, cJ, n), c11);
  ORC_CATCH
}
int orc_differentiate(const double* a11, double* interp9) {
  ORC_TRY
  differentiate(rvpFrom(a11), interp9);
  ORC_CATCH
}
int orc_integrate_interp(const double* interp9, double dt, double* out11) {
  ORC_TRY
  rvpTo(integrateInterp(interp9, dt), out11);
  ORC_CATCH
}
int orc_compensate_jac(const double* calib32, uint32_t flags, const double* raw_gyro3, const double* raw_accel3, double* gyro3, double* accel3,
                       double* cmJ /* 6 x (6 + E) */) {
  ORC_TRY
  const ImuIdx ix(flags);
  compensateAndJac(calibMats(calib32), ix, raw_gyro3, raw_accel3, gyro3, accel3, cmJ);
  ORC_CATCH
}
int orc_imu_boxplus(double* calib32, const double* step, uint32_t flags) {
  ORC_TRY
  imuCalibBoxPlus(calib32, step, flags);
  ORC_CATCH
}
// integrateMeasurements (PreIntegration.cpp:277-306) on raw samples: out = q(4) dV(3) dP(3) dt
int orc_integrate_measurements(int64_t n_samples, const int64_t* sample_time_ns, const float* gyro, const float* accel,
                               const double* calib32, int64_t start_us, int64_t end_us, double* out11) {
  ORC_TRY
  ImuSamples m;
  m.timeNs.assign(sample_time_ns, sample_time_ns + n_samples);
  m.gyro.assign(gyro, gyro + 3 * n_samples);
  m.accel.assign(accel, accel + 3 * n_samples);
  const Rvp r = integrateMeasurements(m, measModelFromCalib(calib32), start_us, end_us);
  for (int i = 0; i < 4; i++) out11[i] = r.q[i];
  for (int i = 0; i < 3; i++) { out11[4 + i] = r.dV[i]; out11[7 + i] = r.dP[i]; }
  out11[10] = r.dt;
  ORC_CATCH
}
int orc_rs_set_gravity_var(OrcHandle* h, int store, int gravity_var) {
  ORC_TRY
  for (RsSource& src : h->rs) if (src.store == store) src.gravityVar = gravity_var;
  ORC_CATCH
}
int orc_rs_refresh(OrcHandle* h, int store) {
  ORC_TRY
  rsRefresh(h, store);
  ORC_CATCH
}
int orc_get_store_aux(OrcHandle* h, int store, double* out, int64_t n) {
  const auto& a = h->eng.P.stores[store].aux;
  if ((int64_t)a.size() != n) return 1;
  std::memcpy(out, a.data(), n * 8);
  return 0;
}

// one accepted-path LM iteration, wall-clock split in ms: {gradhess, damp, factor, solve, apply, cost, total}
// CPU-baseline mode (bench.py): `threads` OpenMP threads for the factor loops and landmark elimination;
// `blas_path` (may be null) = a shared library exporting LAPACK dpbtrf/dpbtrs under `prefix` (scipy's
// OpenBLAS uses "scipy_"), used for the reduced-system Cholesky.  threads <= 1 restores the checker path.
int orc_set_parallel(OrcHandle* h, int threads, const char* blas_path, const char* prefix) {
  ORC_TRY
  h->eng.threads = threads < 1 ? 1 : threads;
  h->sol.pbtrf = nullptr; h->sol.pbtrs = nullptr;
  if (threads > 1 && blas_path) {
    void* lib = dlopen(blas_path, RTLD_NOW | RTLD_GLOBAL);
    if (!lib) throw std::runtime_error(std::string("dlopen failed: ") + dlerror());
    const std::string pre = prefix ? prefix : "";
    h->sol.pbtrf = (SchurSkylineSolver::pbtrf_fn)dlsym(lib, (pre + "dpbtrf_").c_str());
    h->sol.pbtrs = (SchurSkylineSolver::pbtrs_fn)dlsym(lib, (pre + "dpbtrs_").c_str());
    if (!h->sol.pbtrf || !h->sol.pbtrs) throw std::runtime_error("dpbtrf_/dpbtrs_ not found in " + std::string(blas_path));
    typedef void (*setn_fn)(int);
    if (auto setn = (setn_fn)dlsym(lib, (pre + "openblas_set_num_threads").c_str())) setn(threads);
  }
  ORC_CATCH
}

int orc_lm_iteration_timed(OrcHandle* h, double lambda, double* ms, double* costBefore, double* costAfter) {
  ORC_TRY
  using clk = std::chrono::steady_clock;
  auto ms_since = [](clk::time_point a) { return std::chrono::duration<double, std::milli>(clk::now() - a).count(); };
  Engine& e = h->eng;
  if (!h->analyzed) { h->sol.analyze(h->eng); h->analyzed = true; }
  const auto t0 = clk::now();
  if (e.preStep) e.preStep(); // the reference's preStepCallback is inside the iteration (Optimizer.cpp:801-805)
  std::fill(e.grad.begin(), e.grad.end(), 0.0);
  std::fill(e.hess.begin(), e.hess.end(), 0.0);
  const double c0 = e.computeGradHess(e.grad.data(), e.hess.data(), true, false);
  ms[0] = ms_since(t0);
  auto t1 = clk::now();
  e.addDamping(lambda);
  ms[1] = ms_since(t1);
  t1 = clk::now();
  h->sol.factor(e.hess.data());
  ms[2] = ms_since(t1);
  t1 = clk::now();
  h->sol.solveOnly(e.hess.data(), e.grad.data(), e.step.data());
  ms[3] = ms_since(t1);
  t1 = clk::now();
  ms[7] = vdot(e.step, e.grad); // step.dot(grad) before the sign flip (Optimizer.cpp:833)
  for (double& s : e.step) s = -s;
  e.backup();
  double rr[3];
  e.applyStep(e.step.data(), rr);
  ms[4] = ms_since(t1);
  t1 = clk::now();
  CostStats cs;
  const double c1 = e.computeCost(true, &cs);
  ms[5] = ms_since(t1);
  e.restore();
  ms[6] = ms_since(t0);
  if (costBefore) *costBefore = c0;
  if (costAfter) *costAfter = c1;
  ORC_CATCH
}

// Lie / camera unit entry points for tests/test_oracle_math.py
// fn: 0 so3Exp(3->4) 1 so3Log(4->3) 2 se3Exp(6->7) 3 se3Log(7->6) 4 so3LeftJacobian(3->9)
//     5 so3LeftJacobianInverse(3->9) 6 se3LeftJacobianInverse(6->36) 7 se3Adj(7->36) 8 se3Inv(7->7)
//     9 se3Mul(14->7) 10 s2Ortho(3->6)
int orc_lie(int fn, const double* in, double* out) {
  switch (fn) {
    case 0: so3Exp(in, out); break;
    case 1: so3Log(in, out); break;
    case 2: se3Exp(in, out); break;
    case 3: se3Log(in, out); break;
    case 4: so3LeftJacobian(in, out); break;
    case 5: so3LeftJacobianInverse(in, out); break;
    case 6: se3LeftJacobianInverse(in, out); break;
    case 7: se3Adj(in, out); break;
    case 8: se3Inv(in, out); break;
    case 9: se3Mul(in, in + 7, out); break;
    case 10: s2Ortho(in, out); break;
    default: return 1;
  }
  return 0;
}
int orc_fisheye624(const double* p, const double* prm, double* uv, double* dPoint, double* dParam) {
  return fisheye624Project(p, prm, uv, dPoint, dParam) ? 1 : 0;
}
int orc_boxplus(int kind, uint32_t flags, double aux, double* x, const double* s, int tdim, double* ratio) {
  const double r = applyBoxPlus(kind, flags, aux, x, s, tdim);
  if (ratio) *ratio = r;
  return 0;
}

} // extern "C"
