// small_thing/Factor.h — factor payloads and factor stores of the host mirror.
//
// The reference binds arbitrary C++ functors (lib/small_thing/Factor.h:96-141); the device path
// evaluates a fixed catalogue instead — every FactorStore instantiation reachable from ark_vi_ba
//  — so the payload types below carry only the DATA each of those functors holds,
// under the functor's own name where it has one (viba/problem/VisualFactor.cpp:84-120,122-214,
// InertialFactor.cpp:126,190-301) and a descriptive name for the lambdas of RandomWalkFactor.cpp /
// PriorFactor.cpp / OmegaPriorFactor.cpp.  A type that is not in the catalogue does not compile
// against Optimizer::addFactor: there is no silent host path.
//
// Each payload type F provides:
//   static const char* name();                         // FactorStoreBase::name()
//   static int32_t abiType(int nvars);                 // VB_F_* for this arity, -1 if unsupported
//   static const int32_t* kinds(int nvars);            // expected VB_VAR_* per argument
//   static constexpr int kConsts;  void consts(double*) const;   // per-factor constants
//   static constexpr bool kHasPrecision;               // takes a precision matrix in addFactor
//   optional: static void storeAux(const std::vector<F>&, std::vector<double>&);  // store-level data
#pragma once

#include <array>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "../viba_b200.h"
#include "SoftLoss.h"
#include "Variable.h"

namespace small_thing {

// lib/motion/preintegration/PreIntegration.h: what an inertial factor reads through its reference
// (InertialFactor.cpp:126).  250 doubles, the record layout of the C ABI's store aux.
struct PreIntegrationData {
  std::array<double, 4> R{{0, 0, 0, 1}};  // quaternion x,y,z,w of the preintegrated rotation
  Vec3 dV{}, dP{};
  double dtSec = 0.0;
  std::array<double, 9 * 23> J{};         // d(rot,vel,pos)/d(calibration tangent), row-major
  std::array<double, 32> evalPoint{};     // calibration the integral was computed at
};
static_assert(sizeof(PreIntegrationData) == 250 * sizeof(double), "PreIntegrationData must be 250 packed doubles");

// lib/motion/preintegration/RollingShutterData.h: per-rig table of mid->t motion integrals and the
// interpolation data between them (RollingShutterData.cpp:16-65), as the device reads it:
// table[0] = n ; n x (q(4), dV(3), dP(3), dtSec) from table[1] ; (n-1) x (gyro, accel, dPcorr) from
// table[1 + 32*11].
struct RollingShutterData {
  static constexpr int kMaxSamples = 32;
  static constexpr int kStride = 1 + kMaxSamples * 11 + (kMaxSamples - 1) * 9;
  Vec3 gravityWorld{{0.0, 0.0, -9.81}};
  std::array<double, kStride> table{};
};

namespace detail {
template <typename F, typename = void>
struct HasStoreAux : std::false_type {};
template <typename F>
struct HasStoreAux<F, std::void_t<decltype(F::storeAux(std::declval<const std::vector<F>&>(), std::declval<std::vector<double>&>()))>>
    : std::true_type {};

inline void appendPreInt(const PreIntegrationData* p, std::vector<double>& out) {
  if (!p) throw std::runtime_error("small_thing: inertial factor without preintegration data");
  const double* d = reinterpret_cast<const double*>(p);
  out.insert(out.end(), d, d + 250);
}
}  // namespace detail

// ---- B1: viba/problem/VisualFactor.cpp:36-120 -----------------------------------------------
struct VisualFactor {
  Vec2 projBaseRes{};                              // observation in pixels
  std::array<double, 4> sqrtH_BaseRes{{1, 0, 0, 1}};  // row-major 2x2
  static const char* name() { return "VisualFactor"; }
  static int32_t abiType(int n) { return n == 4 ? VB_F_VISUAL : -1; }
  static const int32_t* kinds(int) {
    static const int32_t k[] = {VB_VAR_VEC3, VB_VAR_SE3, VB_VAR_SE3, VB_VAR_CAM};
    return k;
  }
  static constexpr int kConsts = 6;
  static constexpr bool kHasPrecision = false;
  void consts(double* o) const {
    o[0] = projBaseRes[0]; o[1] = projBaseRes[1];
    for (int i = 0; i < 4; i++) o[2 + i] = sqrtH_BaseRes[i];
  }
};

// ---- B2: VisualFactor.cpp:122-214 (adds the world velocity of the rig) ----------------------
struct RollingShutterVisualFactor {
  Vec2 projBaseRes{};
  std::array<double, 4> sqrtH_BaseRes{{1, 0, 0, 1}};
  int64_t rigIndex = 0;                     // row of the per-rig table
  double imageHeight = 0.0;
  const RollingShutterData* rsData = nullptr;  // user-owned, refreshed between iterations (VisualFactor.cpp:213)
  static const char* name() { return "RollingShutterVisualFactor"; }
  static int32_t abiType(int n) { return n == 5 ? VB_F_VISUAL_RS : -1; }
  static const int32_t* kinds(int) {
    static const int32_t k[] = {VB_VAR_VEC3, VB_VAR_SE3, VB_VAR_SE3, VB_VAR_CAM, VB_VAR_VEC3};
    return k;
  }
  static constexpr int kConsts = 8;
  static constexpr bool kHasPrecision = false;
  void consts(double* o) const {
    o[0] = projBaseRes[0]; o[1] = projBaseRes[1];
    for (int i = 0; i < 4; i++) o[2 + i] = sqrtH_BaseRes[i];
    o[6] = (double)rigIndex;
    o[7] = imageHeight;
  }
  // store aux = gravity(3), number of rigs, then one table per rig index
  static void storeAux(const std::vector<RollingShutterVisualFactor>& fs, std::vector<double>& out) {
    int64_t nRigs = 0;
    for (const auto& f : fs) nRigs = f.rigIndex + 1 > nRigs ? f.rigIndex + 1 : nRigs;
    out.assign(4 + (size_t)nRigs * RollingShutterData::kStride, 0.0);
    out[3] = (double)nRigs;
    for (const auto& f : fs) {
      if (!f.rsData || f.rigIndex < 0) throw std::runtime_error("small_thing: RS factor without rolling-shutter data");
      for (int i = 0; i < 3; i++) out[i] = f.rsData->gravityWorld[i];
      double* dst = out.data() + 4 + (size_t)f.rigIndex * RollingShutterData::kStride;
      for (int i = 0; i < RollingShutterData::kStride; i++) dst[i] = f.rsData->table[i];
    }
  }
};

// ---- B4: InertialFactor.cpp:19-127 (precision = inverse rvp covariance, fixed at creation :313) ----
struct InertialFactor {
  const PreIntegrationData* preInt = nullptr;
  static const char* name() { return "InertialFactor"; }
  static int32_t abiType(int n) { return n == 6 ? VB_F_INERTIAL : -1; }
  static const int32_t* kinds(int) {
    static const int32_t k[] = {VB_VAR_IMUCALIB, VB_VAR_SE3, VB_VAR_VEC3, VB_VAR_SE3, VB_VAR_VEC3, VB_VAR_S2};
    return k;
  }
  static constexpr int kConsts = 0;  // the 9x9 precision matrix is the payload
  static constexpr bool kHasPrecision = true;
  static constexpr int kPrecisionDim = 9;
  void consts(double*) const {}
  static void storeAux(const std::vector<InertialFactor>& fs, std::vector<double>& out) {
    for (const auto& f : fs) detail::appendPreInt(f.preInt, out);
  }
};

// ---- B5/B6: InertialFactor.cpp:190-301; 9 arguments = both ends share the IMU extrinsics,
// 10 arguments = window boundary with separate extrinsics ----
struct SecondaryImuInertialFactor {
  const PreIntegrationData* preInt = nullptr;
  static const char* name() { return "SecondaryImuInertialFactor"; }
  static int32_t abiType(int n) { return n == 9 ? VB_F_INERTIAL_SEC : n == 10 ? VB_F_INERTIAL_SEC_SPLIT : -1; }
  static const int32_t* kinds(int n) {
    static const int32_t k9[] = {VB_VAR_IMUCALIB, VB_VAR_SE3, VB_VAR_VEC3, VB_VAR_VEC3, VB_VAR_SE3,
                                 VB_VAR_VEC3, VB_VAR_VEC3, VB_VAR_SE3, VB_VAR_S2};
    static const int32_t k10[] = {VB_VAR_IMUCALIB, VB_VAR_SE3, VB_VAR_VEC3, VB_VAR_VEC3, VB_VAR_SE3,
                                  VB_VAR_SE3, VB_VAR_VEC3, VB_VAR_VEC3, VB_VAR_SE3, VB_VAR_S2};
    return n == 9 ? k9 : k10;
  }
  static constexpr int kConsts = 0;
  static constexpr bool kHasPrecision = true;
  static constexpr int kPrecisionDim = 9;
  void consts(double*) const {}
  static void storeAux(const std::vector<SecondaryImuInertialFactor>& fs, std::vector<double>& out) {
    for (const auto& f : fs) detail::appendPreInt(f.preInt, out);
  }
};

// ---- B7-B10: RandomWalkFactor.cpp:16-168, r = D .* boxMinus(next, prev) ----
template <int N, int32_t ABI, int32_t KIND>
struct RandomWalkFactorT {
  std::array<double, N> diagSqrtH{};
  static const char* name() { return ABI == VB_F_RW_IMUCALIB ? "ImuCalibRandomWalkFactor" : ABI == VB_F_RW_CAM ? "CamIntrRandomWalkFactor" : "SE3RandomWalkFactor"; }
  static int32_t abiType(int n) { return n == 2 ? ABI : -1; }
  static const int32_t* kinds(int) {
    static const int32_t k[] = {KIND, KIND};
    return k;
  }
  static constexpr int kConsts = N;
  static constexpr bool kHasPrecision = false;
  void consts(double* o) const { for (int i = 0; i < N; i++) o[i] = diagSqrtH[i]; }
};
using ImuCalibRandomWalkFactor = RandomWalkFactorT<23, VB_F_RW_IMUCALIB, VB_VAR_IMUCALIB>;
using CamIntrRandomWalkFactor = RandomWalkFactorT<17, VB_F_RW_CAM, VB_VAR_CAM>;
using SE3RandomWalkFactor = RandomWalkFactorT<6, VB_F_RW_SE3, VB_VAR_SE3>;

// ---- B11-B14: PriorFactor.cpp:34-176 ----
struct ImuCalibPriorFactor {   // r = D .* boxMinus(x, prior)
  std::array<double, 23> diagSqrtH{};
  std::array<double, 32> prior{};
  static const char* name() { return "ImuCalibPriorFactor"; }
  static int32_t abiType(int n) { return n == 1 ? VB_F_PRIOR_IMUCALIB : -1; }
  static const int32_t* kinds(int) { static const int32_t k[] = {VB_VAR_IMUCALIB}; return k; }
  static constexpr int kConsts = 55;
  static constexpr bool kHasPrecision = false;
  void consts(double* o) const {
    for (int i = 0; i < 23; i++) o[i] = diagSqrtH[i];
    for (int i = 0; i < 32; i++) o[23 + i] = prior[i];
  }
};

struct CamIntrPriorFactor {
  std::array<double, 17> diagSqrtH{};
  std::array<double, 17> prior{};   // 15 parameters, readout time, time offset
  static const char* name() { return "CamIntrPriorFactor"; }
  static int32_t abiType(int n) { return n == 1 ? VB_F_PRIOR_CAM : -1; }
  static const int32_t* kinds(int) { static const int32_t k[] = {VB_VAR_CAM}; return k; }
  static constexpr int kConsts = 34;
  static constexpr bool kHasPrecision = false;
  void consts(double* o) const {
    for (int i = 0; i < 17; i++) { o[i] = diagSqrtH[i]; o[17 + i] = prior[i]; }
  }
};

struct SE3PriorFactor {          // r = D .* log(T * prior^-1)  (PriorFactor.cpp:41-51,148-155)
  std::array<double, 6> diagSqrtH{};
  SE3 priorInverse;              // prior^-1, so that the device forms T * priorInverse directly
  static SE3 inverse(const SE3& T) {
    // R(q)^T t with the quaternion conjugate; t' = -R^T t
    const double x = -T.q[0], y = -T.q[1], z = -T.q[2], w = T.q[3];
    const double tx = 2.0 * (y * T.t[2] - z * T.t[1]), ty = 2.0 * (z * T.t[0] - x * T.t[2]), tz = 2.0 * (x * T.t[1] - y * T.t[0]);
    SE3 r;
    r.q = {{x, y, z, w}};
    r.t = {{-(T.t[0] + w * tx + (y * tz - z * ty)), -(T.t[1] + w * ty + (z * tx - x * tz)), -(T.t[2] + w * tz + (x * ty - y * tx))}};
    return r;
  }
  static const char* name() { return "SE3PriorFactor"; }
  static int32_t abiType(int n) { return n == 1 ? VB_F_PRIOR_SE3 : -1; }
  static const int32_t* kinds(int) { static const int32_t k[] = {VB_VAR_SE3}; return k; }
  static constexpr int kConsts = 13;
  static constexpr bool kHasPrecision = false;
  void consts(double* o) const {
    for (int i = 0; i < 6; i++) o[i] = diagSqrtH[i];
    for (int i = 0; i < 4; i++) o[6 + i] = priorInverse.q[i];
    for (int i = 0; i < 3; i++) o[10 + i] = priorInverse.t[i];
  }
};

// ---- B15/B16: OmegaPriorFactor.cpp:24-53, sigma = 10 deg/s (Constants.h:19) on the device ----
struct OmegaPriorFactor {
  Vec3 omegaRadSec{};
  static const char* name() { return "OmegaPriorFactor"; }
  static int32_t abiType(int n) { return n == 1 ? VB_F_OMEGA_PRIOR : n == 2 ? VB_F_OMEGA_PRIOR_SEC : -1; }
  static const int32_t* kinds(int n) {
    static const int32_t k1[] = {VB_VAR_VEC3};
    static const int32_t k2[] = {VB_VAR_VEC3, VB_VAR_SE3};   // body omega, IMU extrinsics
    return n == 1 ? k1 : k2;
  }
  static constexpr int kConsts = 3;
  static constexpr bool kHasPrecision = false;
  void consts(double* o) const { for (int i = 0; i < 3; i++) o[i] = omegaRadSec[i]; }
};

// ---- stores -----------------------------------------------------------------------------------
// Factor.h:33-90: what the optimizer and the introspection users (Histograms.cpp, Stats.cpp) see
struct FactorStoreBase {
  virtual ~FactorStoreBase() = default;
  virtual std::string name() const = 0;
  virtual int64_t numCosts() const = 0;
  virtual int numVars() const = 0;
  virtual VarBase* costVar(int64_t k, int i) const = 0;
  virtual int32_t abiType() const = 0;
  virtual int constsPerFactor() const = 0;
  virtual void fillConsts(double* out) const = 0;          // [numCosts * constsPerFactor]
  virtual void fillAux(std::vector<double>& out) const = 0;
  const Loss* loss = nullptr;                                // pointer, read when the graph is exported
};

template <typename F>
struct FactorStore : FactorStoreBase {
  explicit FactorStore(int nvars_) : nvars(nvars_) {}
  std::string name() const override { return F::name(); }
  int64_t numCosts() const override { return (int64_t)factors.size(); }
  int numVars() const override { return nvars; }
  VarBase* costVar(int64_t k, int i) const override { return vars[(size_t)k * nvars + i]; }
  int32_t abiType() const override { return F::abiType(nvars); }
  int constsPerFactor() const override { return F::kConsts + (F::kHasPrecision ? precisionDim * precisionDim : 0); }
  void fillConsts(double* out) const override {
    const int cpf = constsPerFactor();
    for (size_t k = 0; k < factors.size(); k++) {
      factors[k].consts(out + k * cpf);
      if (F::kHasPrecision)
        for (int i = 0; i < precisionDim * precisionDim; i++) out[k * cpf + F::kConsts + i] = precision[k * precisionDim * precisionDim + i];
    }
  }
  void fillAux(std::vector<double>& out) const override {
    out.clear();
    if constexpr (detail::HasStoreAux<F>::value) F::storeAux(factors, out);
  }

  int nvars;
  int precisionDim = 0;
  std::vector<F> factors;           // boundFactors of the reference, payload only
  std::vector<VarBase*> vars;       // raw pointers to user-owned variables (Factor.h:123)
  std::vector<double> precision;    // row-major, per factor
};

}  // namespace small_thing
