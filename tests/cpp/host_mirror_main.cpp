// Test driver for the C++17 host mirror (include/small_thing/*.h): reads a session written by
// tests/test_host_mirror.py, rebuilds it through small_thing::Optimizer exactly as a user of the
// reference would (Variables owned here, one addFactor call per factor), and writes back either
//   * the graph the mirror hands to the C ABI (host only; compared array by array with the Python
//     builder's graph), or
//   * with --optimize N: the Summary and the optimised values (needs a B200; throws otherwise).
//
// File format, little endian, arrays back to back:
//   int64 nv | int32 kind[nv] | int32 tdim[nv] | uint32 flags[nv] | double aux[nv] | uint8 const[nv]
//   int64 dataLen | double data[dataLen] | int64 np | int32 pointIds[np] | int32 nstores
//   per store: int32 type, nvars | int64 count | int32 cpf | double lossA, lossK | int64 auxLen
//              | int32 vars[count*nvars] | double consts[count*cpf] | double aux[auxLen]
#include <small_thing/Optimizer.h>
#include <baspacho/baspacho/Solver.h>
#include <baspacho/baspacho/Utils.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <deque>
#include <fstream>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

using namespace small_thing;

namespace {

struct Reader {
  std::ifstream f;
  explicit Reader(const char* p) : f(p, std::ios::binary) {
    if (!f) throw std::runtime_error(std::string("cannot open ") + p);
  }
  template <typename T>
  T one() {
    T v;
    f.read(reinterpret_cast<char*>(&v), sizeof(T));
    return v;
  }
  template <typename T>
  std::vector<T> many(int64_t n) {
    std::vector<T> v((size_t)n);
    f.read(reinterpret_cast<char*>(v.data()), (std::streamsize)(sizeof(T) * (size_t)n));
    if (!f) throw std::runtime_error("short read");
    return v;
  }
};

struct Writer {
  std::ofstream f;
  explicit Writer(const char* p) : f(p, std::ios::binary) {}
  template <typename T>
  void one(T v) { f.write(reinterpret_cast<const char*>(&v), sizeof(T)); }
  template <typename T>
  void many(const T* p, size_t n) { f.write(reinterpret_cast<const char*>(p), (std::streamsize)(sizeof(T) * n)); }
  template <typename T>
  void many(const std::vector<T>& v) { many(v.data(), v.size()); }
};

template <typename T>
Variable<T>& as(std::vector<std::unique_ptr<VarBase>>& vars, int32_t id) {
  auto* p = dynamic_cast<Variable<T>*>(vars.at((size_t)id).get());
  if (!p) throw std::runtime_error("variable " + std::to_string(id) + " has an unexpected type");
  return *p;
}

template <int N>
std::array<double, N> arr(const double* p) {
  std::array<double, N> a;
  for (int i = 0; i < N; i++) a[i] = p[i];
  return a;
}

}  // namespace

// host-side contract of the mirror, no device: every misuse the reference rejects with BASPACHO_CHECK / a
// compile error is a std::runtime_error here
template <typename Fn>
static bool throwsRuntimeError(Fn&& fn) {
  try {
    fn();
  } catch (const std::runtime_error&) {
    return true;
  }
  return false;
}

static int selftest() {
  int bad = 0;
  auto expect = [&](bool ok, const char* what) {
    if (!ok) {
      std::fprintf(stderr, "selftest failed: %s\n", what);
      bad++;
    }
  };
  Variable<Vec3> X, X2, vel;
  Variable<SE3> T_bw, T_cb;
  Variable<CameraModelParam> cam;
  Variable<ImuCalibParam> imu;
  Variable<S2> g;
  HuberLossWithCutoff loss(1.0, 3.0), otherLoss(2.0, 3.0);
  const VisualFactor vf{{{10.0, 20.0}}, {{0.7, 0.0, 0.0, 0.7}}};
  {
    Optimizer opt;
    auto [store, k] = opt.addFactor(vf, loss, X, T_bw, T_cb, cam);
    expect(k == 0 && store->numCosts() == 1 && store->name() == "VisualFactor" && store->costVar(0, 1) == &T_bw, "store introspection");
    expect(opt.addFactor(vf, loss, X2, T_bw, T_cb, cam).second == 1, "second factor lands in the same store");
    expect(opt.factorStores.size() == 1, "one store per factor type");
    expect(throwsRuntimeError([&] { opt.addFactor(vf, otherLoss, X, T_bw, T_cb, cam); }), "two different losses in one store");
    expect(throwsRuntimeError([&] { opt.addFactor(vf, loss, X, T_bw, T_cb); }), "wrong arity");
    expect(throwsRuntimeError([&] { opt.addFactor(vf, loss, T_bw, X, T_cb, cam); }), "wrong variable types");
    const double info[81] = {};
    expect(throwsRuntimeError([&] { opt.addFactor(vf, info, loss, X, T_bw, T_cb, cam); }), "precision matrix on a factor that takes none");
    PreIntegrationData pre;
    Variable<SE3> T2;
    Variable<Vec3> v2;
    expect(throwsRuntimeError([&] { opt.addFactor(InertialFactor{&pre}, loss, imu, T_bw, vel, T2, v2, g); }), "inertial factor without precision matrix");
    // registration: points first, then per store / argument / factor; constants keep kConstantVar
    g.setConstant();
    opt.registerVariable(X);
    opt.registerVariable(X2);
    opt.registeredVariablesToEliminationRange();
    opt.registerAllVariables();
    expect(X.index == 0 && X2.index == 1 && T_bw.index == 2 && T_cb.index == 3 && cam.index == 4, "registration order");
    expect(opt.elimRanges.size() == 2 && opt.elimRanges[1] == 2, "elimination range");
    expect(opt.totalVarDimensionality() == 3 + 3 + 6 + 6 + 15 && opt.totalErrorDimensionality() == 4, "dimensionalities");
    expect(throwsRuntimeError([&] { X.setConstant(); }), "setConstant on a registered variable");
    expect(g.isSetToConstant() && !g.isRegistered(), "constant variable");
    auto ex = opt.exportGraph();
    expect(ex->desc.num_vars == 5 && ex->desc.num_points == 2 && ex->stores[0].count == 2 && ex->stores[0].loss_a == 1.0 && ex->stores[0].loss_k == 3.0, "exported graph");
    loss.setSize(1.5, 4.0);   // the loss is held by pointer: later changes are what the device sees (Factor.h:141)
    expect(opt.exportGraph()->stores[0].loss_a == 1.5, "loss read at export time");
    opt.unregisterAllVariables();
    expect(X.isUnset() && cam.isUnset() && opt.paramSizes.empty() && g.isSetToConstant(), "unregisterAllVariables");
  }
  {
    Optimizer opt;   // the elimination range must hold 3-D points only, and there is one range
    opt.registerVariable(T_bw);
    opt.registeredVariablesToEliminationRange();
    opt.addFactor(vf, X, T_bw, T_cb, cam);
    expect(throwsRuntimeError([&] { opt.exportGraph(); }), "pose inside the elimination range");
    opt.unregisterAllVariables();
  }
  expect(throwsRuntimeError([] { HuberLossWithCutoff l(0.0, 1.0); }) && throwsRuntimeError([] { HuberLossWithCutoff l(2.0, 1.0); }), "loss parameter checks");
  expect(HuberLossWithCutoff(1.0, 3.0).val(0.5) == 0.5 && HuberLossWithCutoff(1.0, 3.0).val(4.0) == 3.0 && HuberLossWithCutoff(1.0, 3.0).val(100.0) == 5.0, "loss values");
  std::printf("selftest: %s\n", bad ? "FAILED" : "ok");
  return bad ? 1 : 0;
}

int run(int argc, char** argv) {
  if (argc >= 2 && std::string(argv[1]) == "--selftest") return selftest();
  if (argc < 3) {
    std::fprintf(stderr, "usage: %s in.bin out.bin [--optimize N]\n", argv[0]);
    return 2;
  }
  int optimizeIters = -1;
  if (argc >= 5 && std::string(argv[3]) == "--optimize") optimizeIters = std::atoi(argv[4]);
  const bool withCallback = argc >= 6 && std::string(argv[5]) == "--callback";
  const bool withPcg = argc >= 6 && std::string(argv[5]) == "--pcg-jacobi";   // Settings::solverType = Solver_PCG_Jacobi
  const bool withGs = argc >= 6 && std::string(argv[5]) == "--pcg-gauss-seidel";
  const bool withFacade = argc >= 4 && std::string(argv[3]) == "--facade";
  const bool withCovariance = argc >= 4 && std::string(argv[3]) == "--covariance";

  Reader in(argv[1]);
  const int64_t nv = in.one<int64_t>();
  auto kind = in.many<int32_t>(nv);
  auto tdim = in.many<int32_t>(nv);
  auto flags = in.many<uint32_t>(nv);
  auto aux = in.many<double>(nv);
  auto isConst = in.many<uint8_t>(nv);
  const int64_t dataLen = in.one<int64_t>();
  auto data = in.many<double>(dataLen);

  // user-owned variables, created in the session's own order
  std::vector<std::unique_ptr<VarBase>> vars;
  const double* d = data.data();
  for (int64_t i = 0; i < nv; i++) {
    std::unique_ptr<VarBase> v;
    switch (kind[i]) {
      case VB_VAR_VEC3: v = std::make_unique<Variable<Vec3>>(); break;
      case VB_VAR_VEC2: v = std::make_unique<Variable<Vec2>>(); break;
      case VB_VAR_SE3: v = std::make_unique<Variable<SE3>>(); break;
      case VB_VAR_S2: {
        auto s = std::make_unique<Variable<S2>>();
        s->value.radius = aux[i];
        v = std::move(s);
        break;
      }
      case VB_VAR_CAM: {
        auto c = std::make_unique<Variable<CameraModelParam>>();
        c->value.estimateReadoutTime = flags[i] & VB_CAM_EST_READOUT;
        c->value.estimateTimeOffset = flags[i] & VB_CAM_EST_TIMEOFFSET;
        c->value.hasReadoutTime = flags[i] & VB_CAM_HAS_READOUT;
        v = std::move(c);
        break;
      }
      case VB_VAR_IMUCALIB: {
        auto c = std::make_unique<Variable<ImuCalibParam>>();
        c->value.options = flags[i];
        c->value.tangentDim = tdim[i];
        v = std::move(c);
        break;
      }
      default: throw std::runtime_error("unknown variable kind");
    }
    v->setData(d);
    d += v->getDataDim();
    if (isConst[i]) v->setConstant();
    if (v->getTangentDim() != tdim[i]) throw std::runtime_error("tangent dim mismatch for variable " + std::to_string(i));
    vars.push_back(std::move(v));
  }

  Optimizer opt;
  // registerPointVariables + registeredVariablesToEliminationRange (SingleSessionProblem.cpp:41-45,
  // main_AriaKit_ViBa.cpp:76-77)
  const int64_t np = in.one<int64_t>();
  auto pointIds = in.many<int32_t>(np);
  for (int32_t id : pointIds) opt.registerVariable(as<Vec3>(vars, id));
  opt.registeredVariablesToEliminationRange();

  std::deque<HuberLossWithCutoff> losses;          // addresses must stay valid (Factor.h:141)
  std::deque<std::vector<PreIntegrationData>> preints;
  std::deque<std::vector<RollingShutterData>> rsTables;
  auto lossFor = [&](double a, double k) -> const HuberLossWithCutoff& {
    for (auto& l : losses)
      if (l.a == a && l.k == k) return l;
    losses.emplace_back(a, k);
    return losses.back();
  };

  const int32_t ns = in.one<int32_t>();
  for (int32_t s = 0; s < ns; s++) {
    const int32_t type = in.one<int32_t>(), nvars = in.one<int32_t>();
    const int64_t count = in.one<int64_t>();
    const int32_t cpf = in.one<int32_t>();
    const double lossA = in.one<double>(), lossK = in.one<double>();
    const int64_t auxLen = in.one<int64_t>();
    auto sv = in.many<int32_t>(count * nvars);
    auto sc = in.many<double>(count * cpf);
    auto sa = in.many<double>(auxLen);
    const bool soft = lossA > 0.0;
    if (type == VB_F_INERTIAL || type == VB_F_INERTIAL_SEC || type == VB_F_INERTIAL_SEC_SPLIT) {
      preints.emplace_back((size_t)count);
      std::memcpy(static_cast<void*>(preints.back().data()), sa.data(), sizeof(double) * 250 * (size_t)count);
    }
    if (type == VB_F_VISUAL_RS) {
      const int64_t nRigs = (int64_t)sa[3];
      rsTables.emplace_back((size_t)nRigs);
      for (int64_t r = 0; r < nRigs; r++) {
        auto& t = rsTables.back()[(size_t)r];
        t.gravityWorld = {{sa[0], sa[1], sa[2]}};
        std::memcpy(t.table.data(), sa.data() + 4 + r * RollingShutterData::kStride, sizeof(double) * RollingShutterData::kStride);
      }
    }
    for (int64_t k = 0; k < count; k++) {
      const int32_t* v = sv.data() + k * nvars;
      const double* c = sc.data() + k * cpf;
      switch (type) {
        case VB_F_VISUAL: {
          VisualFactor f{{{c[0], c[1]}}, arr<4>(c + 2)};
          if (soft) opt.addFactor(f, lossFor(lossA, lossK), as<Vec3>(vars, v[0]), as<SE3>(vars, v[1]), as<SE3>(vars, v[2]), as<CameraModelParam>(vars, v[3]));
          else opt.addFactor(f, as<Vec3>(vars, v[0]), as<SE3>(vars, v[1]), as<SE3>(vars, v[2]), as<CameraModelParam>(vars, v[3]));
          break;
        }
        case VB_F_VISUAL_RS: {
          RollingShutterVisualFactor f;
          f.projBaseRes = {{c[0], c[1]}};
          f.sqrtH_BaseRes = arr<4>(c + 2);
          f.rigIndex = (int64_t)c[6];
          f.imageHeight = c[7];
          f.rsData = &rsTables.back()[(size_t)f.rigIndex];
          if (!soft) throw std::runtime_error("RS store without a loss is not part of this test");
          opt.addFactor(f, lossFor(lossA, lossK), as<Vec3>(vars, v[0]), as<SE3>(vars, v[1]), as<SE3>(vars, v[2]),
                        as<CameraModelParam>(vars, v[3]), as<Vec3>(vars, v[4]));
          break;
        }
        case VB_F_INERTIAL: {
          InertialFactor f{&preints.back()[(size_t)k]};
          opt.addFactor(f, c, lossFor(lossA, lossK), as<ImuCalibParam>(vars, v[0]), as<SE3>(vars, v[1]), as<Vec3>(vars, v[2]),
                        as<SE3>(vars, v[3]), as<Vec3>(vars, v[4]), as<S2>(vars, v[5]));
          break;
        }
        case VB_F_INERTIAL_SEC: {
          SecondaryImuInertialFactor f{&preints.back()[(size_t)k]};
          opt.addFactor(f, c, lossFor(lossA, lossK), as<ImuCalibParam>(vars, v[0]), as<SE3>(vars, v[1]), as<Vec3>(vars, v[2]),
                        as<Vec3>(vars, v[3]), as<SE3>(vars, v[4]), as<Vec3>(vars, v[5]), as<Vec3>(vars, v[6]), as<SE3>(vars, v[7]),
                        as<S2>(vars, v[8]));
          break;
        }
        case VB_F_INERTIAL_SEC_SPLIT: {
          SecondaryImuInertialFactor f{&preints.back()[(size_t)k]};
          opt.addFactor(f, c, lossFor(lossA, lossK), as<ImuCalibParam>(vars, v[0]), as<SE3>(vars, v[1]), as<Vec3>(vars, v[2]),
                        as<Vec3>(vars, v[3]), as<SE3>(vars, v[4]), as<SE3>(vars, v[5]), as<Vec3>(vars, v[6]), as<Vec3>(vars, v[7]),
                        as<SE3>(vars, v[8]), as<S2>(vars, v[9]));
          break;
        }
        case VB_F_RW_IMUCALIB:
          opt.addFactor(ImuCalibRandomWalkFactor{arr<23>(c)}, as<ImuCalibParam>(vars, v[0]), as<ImuCalibParam>(vars, v[1]));
          break;
        case VB_F_RW_CAM:
          opt.addFactor(CamIntrRandomWalkFactor{arr<17>(c)}, as<CameraModelParam>(vars, v[0]), as<CameraModelParam>(vars, v[1]));
          break;
        case VB_F_RW_SE3:
          opt.addFactor(SE3RandomWalkFactor{arr<6>(c)}, as<SE3>(vars, v[0]), as<SE3>(vars, v[1]));
          break;
        case VB_F_PRIOR_IMUCALIB:
          opt.addFactor(ImuCalibPriorFactor{arr<23>(c), arr<32>(c + 23)}, as<ImuCalibParam>(vars, v[0]));
          break;
        case VB_F_PRIOR_CAM:
          opt.addFactor(CamIntrPriorFactor{arr<17>(c), arr<17>(c + 17)}, as<CameraModelParam>(vars, v[0]));
          break;
        case VB_F_PRIOR_SE3: {
          SE3PriorFactor f;
          f.diagSqrtH = arr<6>(c);
          f.priorInverse.q = arr<4>(c + 6);
          f.priorInverse.t = arr<3>(c + 10);
          {  // the helper a user calls with the prior itself: inverting twice must give the payload back
            const SE3 back = SE3PriorFactor::inverse(SE3PriorFactor::inverse(f.priorInverse));
            for (int i = 0; i < 4; i++)
              if (std::abs(back.q[i] - f.priorInverse.q[i]) > 1e-14) throw std::runtime_error("SE3PriorFactor::inverse: rotation");
            for (int i = 0; i < 3; i++)
              if (std::abs(back.t[i] - f.priorInverse.t[i]) > 1e-13) throw std::runtime_error("SE3PriorFactor::inverse: translation");
          }
          opt.addFactor(f, as<SE3>(vars, v[0]));
          break;
        }
        case VB_F_OMEGA_PRIOR:
          opt.addFactor(OmegaPriorFactor{arr<3>(c)}, as<Vec3>(vars, v[0]));
          break;
        case VB_F_OMEGA_PRIOR_SEC:
          opt.addFactor(OmegaPriorFactor{arr<3>(c)}, as<Vec3>(vars, v[0]), as<SE3>(vars, v[1]));
          break;
        default: throw std::runtime_error("unknown factor type " + std::to_string(type));
      }
    }
  }

  Writer out(argv[2]);
  if (withFacade) {
    // include/baspacho/baspacho/{Solver,Accessor}.h over the mirror's device problem: the calls of the reference's direct
    // path — order / dataSize / accessor().paramStart, paramSize, block, diagBlock / factor / solve
    vb_problem* p = opt.backendProblem();
    BaSpaCho::Solver solver(p);
    const auto& acc = solver.accessor();
    const int64_t n = solver.order(), ds = solver.dataSize();
    double cost = 0;
    if (vb_eval_grad_hess(p, 1, 1, 0, &cost) != VB_OK || vb_damp(p, 1e-5) != VB_OK) throw std::runtime_error(vb_last_error());
    std::vector<double> hess((size_t)ds), grad((size_t)n), step((size_t)n), ref((size_t)n);
    if (vb_download_hess(p, hess.data(), ds) != VB_OK || vb_download_grad(p, grad.data(), n) != VB_OK) throw std::runtime_error(vb_last_error());
    // accessor: every registered variable's segment has its tangent size, diagonal blocks are symmetric and positive
    int64_t checked = 0, total = 0;
    for (auto& v : vars) {
      if (v->index < 0) continue;
      BASPACHO_CHECK_EQ(acc.paramSize(v->index), (int64_t)v->getTangentDim());
      BASPACHO_CHECK_LE(acc.paramStart(v->index) + acc.paramSize(v->index), n);
      total += acc.paramSize(v->index);
      auto D = acc.diagBlock(hess.data(), v->index);
      for (int64_t a = 0; a < D.rows; a++) { BASPACHO_CHECK_GT(D(a, a), 0.0); checked++; }
    }
    BASPACHO_CHECK_EQ(total, n);
    step = grad;
    std::vector<double> factored = hess;
    solver.factor(factored.data());
    solver.solve(factored.data(), step.data(), n, 1);
    // the same through the plain ABI
    if (vb_upload_hess(p, hess.data(), ds) != VB_OK || vb_upload_grad(p, grad.data(), n) != VB_OK || vb_step_from_grad(p) != VB_OK ||
        vb_factor(p) != VB_OK || vb_solve(p, 0) != VB_OK || vb_download_step(p, ref.data(), n) != VB_OK)
      throw std::runtime_error(vb_last_error());
    double worst = 0, scale = 0, resid = 0;
    for (int64_t i = 0; i < n; i++) { worst = std::max(worst, std::fabs(step[(size_t)i] - ref[(size_t)i])); scale = std::max(scale, std::fabs(ref[(size_t)i])); }
    (void)resid;
    bool threw = false;
    try { solver.factorUpTo(factored.data(), 1); } catch (const std::runtime_error&) { threw = true; }
    out.one<double>((double)n); out.one<double>((double)ds); out.one<double>((double)checked); out.one<double>(worst); out.one<double>(scale);
    out.one<double>(threw ? 1.0 : 0.0); out.one<double>(cost);
    return 0;
  }
  if (withCovariance) {
    // the recipe of the reference's own test (lib/small_thing/tests/TestOptimizer.cpp:52-83) through the mirror:
    // computeCovariances, computeJointCovariances and computeMarginalProblem of the same variables; offsets agree,
    // marginal information x joint covariance = I, the per-variable covariances are the diagonal blocks
    int64_t nReg = 0;
    for (auto& v : vars) nReg = std::max<int64_t>(nReg, v->index + 1);
    const std::vector<int64_t> idx = {nReg - 1, nReg - 3, nReg - 4};
    Optimizer::Settings st;
    st.damping = 1e-5;
    auto covs = opt.computeCovariances(st, idx);
    auto joint = opt.computeJointCovariances(st, {idx, {idx[1]}});
    auto mProb = opt.computeMarginalProblem(st, idx);
    if (covs.size() != idx.size() || joint.size() != 2) throw std::runtime_error("wrong number of covariance blocks");
    const auto& cov = joint[0].first;
    const auto& covOffs = joint[0].second;
    if (covOffs.size() != mProb.offsets.size()) throw std::runtime_error("offset lists differ in size");
    for (size_t i = 0; i < covOffs.size(); i++)
      if (covOffs[i] != mProb.offsets[i]) throw std::runtime_error("offsets differ");
    const int64_t D = cov.rows;
    if (mProb.H.rows != D || (int64_t)mProb.b.size() != D) throw std::runtime_error("marginal problem of the wrong size");
    double delta2 = 0, diagDiff = 0, diagScale = 0;
    for (int64_t r = 0; r < D; r++)
      for (int64_t c = 0; c < D; c++) {
        double t = r == c ? -1.0 : 0.0;
        for (int64_t k = 0; k < D; k++) t += mProb.H(r, k) * cov(k, c);
        delta2 += t * t;
      }
    for (size_t v = 0; v < idx.size(); v++) {
      const int64_t o = covOffs[v], d = covOffs[v + 1] - o;
      if (covs[v].rows != d) throw std::runtime_error("per-variable covariance of the wrong size");
      for (int64_t r = 0; r < d; r++)
        for (int64_t c = 0; c < d; c++) {
          diagDiff = std::max(diagDiff, std::fabs(covs[v](r, c) - cov(o + r, o + c)));
          diagScale = std::max(diagScale, std::fabs(cov(o + r, o + c)));
        }
    }
    // the second list of the joint call is variable idx[1] alone
    const int64_t o1 = covOffs[1], d1 = covOffs[2] - o1;
    for (int64_t r = 0; r < d1; r++)
      for (int64_t c = 0; c < d1; c++) diagDiff = std::max(diagDiff, std::fabs(joint[1].first(r, c) - cov(o1 + r, o1 + c)));
    bool threw = false;
    try { opt.computeMarginalProblem(st, {idx[0], idx[0]}); } catch (const std::runtime_error&) { threw = true; }
    out.one<double>((double)D); out.one<double>(std::sqrt(delta2)); out.one<double>(diagDiff); out.one<double>(diagScale);
    out.one<double>(mProb.cost); out.one<double>(opt.computeCost()); out.one<double>(threw ? 1.0 : 0.0);
    return 0;
  }
  if (optimizeIters >= 0) {
    Optimizer::Settings st;
    st.maxNumIterations = optimizeIters;
    st.log = nullptr;
    const double cost0 = opt.computeCost();
    int logLines = 0, callbacks = 0;
    double calibSeen = 0.0;
    if (withCallback) {
      // Settings::preStepCallback as interfaces/ark/main_AriaKit_ViBa.cpp:95-101 uses it: user code runs at the top of
      // every iteration, reads the CURRENT calibration from its own Variable and rewrites a preintegration record
      // (here: at iteration 2 record 0 of the first inertial store gets dV[0] += 1e-3 and the calibration of that
      // iteration as evalPoint).  The test drives the oracle the same way and compares.
      st.log = [&](const std::string& line) { logLines += (line.find(" cost: ") != std::string::npos); };
      st.preStepCallback = [&](int it) {
        callbacks++;
        if (it != 2 || preints.empty()) return;
        PreIntegrationData& rec = preints.front()[0];
        rec.dV[0] += 1e-3;
        VarBase* calib = nullptr;
        for (auto& v : vars) if (v->kind() == VB_VAR_IMUCALIB && v->index >= 0) { calib = v.get(); break; }
        if (calib) { calib->getData(rec.evalPoint.data()); calibSeen = rec.evalPoint[6]; }
      };
      Optimizer::Settings bad = st;
      bad.solverType = (Optimizer::SolverType)9;   // "Unknown preconditioner" (Optimizer.cpp:243-245): refused up front
      bool threw = false;
      try { opt.optimize(bad); } catch (const std::runtime_error&) { threw = true; }
      if (!threw || callbacks != 0) throw std::runtime_error("an unknown solver type was accepted");
    }
    if (withPcg) { st.solverType = Optimizer::Solver_PCG_Jacobi; st.pcgMaxIterations = 8000; st.pcgDesiredResidual = 1e-12; }
    if (withGs) { st.solverType = Optimizer::Solver_PCG_GaussSeidel; st.pcgMaxIterations = 8000; st.pcgDesiredResidual = 1e-12; }
    Optimizer::Summary sm = opt.optimize(st);
    const double cost1 = opt.computeCost();
    out.one<double>(sm.initialCost);
    out.one<double>(sm.finalCost);
    out.one<double>((double)sm.numIterations);
    out.one<double>(cost0);
    out.one<double>(cost1);
    if (withCallback) { out.one<double>((double)logLines); out.one<double>((double)callbacks); out.one<double>(calibSeen); }
    std::vector<double> buf(32);
    for (auto& v : vars) {   // the user's own Variables now hold the optimised values
      v->getData(buf.data());
      out.many(buf.data(), (size_t)v->getDataDim());
    }
    return 0;
  }

  // host only: the graph as the mirror hands it to vb_problem_create, plus the symbolic phase on it
  auto g = opt.exportGraph();
  std::unordered_map<const VarBase*, int32_t> poolId;
  for (size_t i = 0; i < g->pool.size(); i++) poolId[g->pool[i]] = (int32_t)i;
  std::vector<int32_t> py2cpp((size_t)nv, -1);
  for (int64_t i = 0; i < nv; i++) {
    auto it = poolId.find(vars[(size_t)i].get());
    if (it != poolId.end()) py2cpp[(size_t)i] = it->second;
  }
  const int64_t n = g->desc.num_vars;
  out.one<int64_t>(n);
  out.many(g->kind);
  out.many(g->tdim);
  out.many(g->flags);
  out.many(g->index);
  out.many(g->off);
  out.many(g->aux);
  out.one<int64_t>(g->desc.var_data_len);
  out.many(g->data);
  out.one<int64_t>(g->desc.num_points);
  out.one<int64_t>(nv);
  out.many(py2cpp);
  out.one<int32_t>(g->desc.num_stores);
  for (const auto& s : g->stores) {
    out.one<int32_t>(s.type);
    out.one<int32_t>(s.nvars);
    out.one<int64_t>(s.count);
    out.one<int32_t>(s.consts_per_factor);
    out.one<double>(s.loss_a);
    out.one<double>(s.loss_k);
    out.one<int64_t>(s.aux_len);
    out.many(s.vars, (size_t)(s.count * s.nvars));
    out.many(s.consts, (size_t)(s.count * s.consts_per_factor));
    out.many(s.aux, (size_t)s.aux_len);
  }
  int64_t order = 0, dataSize = 0, numLumps = 0;
  if (vb_symbolic_only(&g->desc, &order, &dataSize, &numLumps, nullptr, 0) != VB_OK)
    throw std::runtime_error(std::string("vb_symbolic_only: ") + vb_last_error());
  out.one<int64_t>(order);
  out.one<int64_t>(dataSize);
  out.one<int64_t>(numLumps);
  out.one<int64_t>(opt.totalVarDimensionality());
  out.one<int64_t>(opt.totalErrorDimensionality());
  return 0;
}

int main(int argc, char** argv) {
  try {
    return run(argc, argv);
  } catch (const std::runtime_error& e) {   // what a caller of the reference would catch
    std::fprintf(stderr, "runtime_error: %s\n", e.what());
    return 3;
  }
}
