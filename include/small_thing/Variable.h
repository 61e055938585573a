// small_thing/Variable.h — host-side mirror of the reference's variable API above the C ABI of
// viba_b200.h.  Same names and meaning as lib/small_thing/Variable.h:222-330
// (VarBase, Variable<T>, kUnsetIndex / kConstantVar, setConstant, isRegistered) so that code written
// against small_thing::Optimizer keeps compiling; the value types are plain structs because
// Eigen / Sophus are not part of this build (their data layouts are kept: SE3 = unit quaternion
// x,y,z,w followed by the translation, as Sophus::SE3d::data()).
//
// Header-only C++17, no dependencies beyond viba_b200.h.  Nothing here computes residuals: the
// retraction, the factors and the solver run on the GPU behind the C ABI; these classes only hold
// user-owned values and describe them to vb_graph_desc.
#pragma once

#include <array>
#include <cstdint>
#include <stdexcept>

#include "../viba_b200.h"

namespace small_thing {

using Vec2 = std::array<double, 2>;
using Vec3 = std::array<double, 3>;

// Sophus::SE3d layout (Variable.h:97-128 reads/writes it as 7 doubles)
struct SE3 {
  std::array<double, 4> q{{0.0, 0.0, 0.0, 1.0}};  // x, y, z, w
  Vec3 t{{0.0, 0.0, 0.0}};
};

// point on a sphere of fixed radius (Variable.h:160-221); gravity in ark_vi_ba
struct S2 {
  Vec3 vec{{0.0, 0.0, -9.81}};
  double radius = 9.81;
};

// interfaces/ark/camera_model/CameraModelParam.h:21-55,124-149: Fisheye624 parameters plus the
// optional rolling-shutter readout time and camera time offset (tangent dim 15..17)
struct CameraModelParam {
  std::array<double, 15> params{};
  double readoutTimeSec = 0.0;
  double timeOffsetSec = 0.0;
  bool hasReadoutTime = false;
  bool estimateReadoutTime = false;
  bool estimateTimeOffset = false;
};

// interfaces/ark/imu_model/ImuCalibParam.h:23-75: 32 stored doubles, tangent dim <= 23 selected by
// the ImuCalibrationOptions bits
struct ImuCalibParam {
  std::array<double, 32> data{};
  uint32_t options = 255u;
  int tangentDim = 23;
};

constexpr int64_t kUnsetIndex = -1;
constexpr int64_t kConstantVar = -2;

struct VarBase {
  virtual ~VarBase() = default;

  // Variable.h:228-233: constness cannot change once registered
  void setConstant(bool constant = true) {
    if (isRegistered()) throw std::runtime_error("small_thing: setConstant on a registered variable");
    index = constant ? kConstantVar : kUnsetIndex;
  }
  bool isSetToConstant() const { return index == kConstantVar; }
  bool isRegistered() const { return index >= 0; }
  bool isUnset() const { return index == kUnsetIndex; }

  // description for vb_graph_desc
  virtual int32_t kind() const = 0;        // VB_VAR_*
  virtual int getDataDim() const = 0;      // doubles in var_data
  virtual int getTangentDim() const = 0;
  virtual uint32_t flags() const { return 0u; }
  virtual double aux() const { return 0.0; }
  virtual void getData(double* out) const = 0;   // Variable.h:332-349 (backup order == vb layout)
  virtual void setData(const double* in) = 0;

  int64_t index = kUnsetIndex;
};

template <typename DType>
struct VarSpec;

template <>
struct VarSpec<Vec3> {
  static constexpr int32_t kKind = VB_VAR_VEC3;
  static constexpr int kData = 3;
  static int tangentDim(const Vec3&) { return 3; }
  static uint32_t flags(const Vec3&) { return 0u; }
  static double aux(const Vec3&) { return 0.0; }
  static void get(const Vec3& v, double* o) { for (int i = 0; i < 3; i++) o[i] = v[i]; }
  static void set(Vec3& v, const double* in) { for (int i = 0; i < 3; i++) v[i] = in[i]; }
};

template <>
struct VarSpec<Vec2> {
  static constexpr int32_t kKind = VB_VAR_VEC2;
  static constexpr int kData = 2;
  static int tangentDim(const Vec2&) { return 2; }
  static uint32_t flags(const Vec2&) { return 0u; }
  static double aux(const Vec2&) { return 0.0; }
  static void get(const Vec2& v, double* o) { o[0] = v[0]; o[1] = v[1]; }
  static void set(Vec2& v, const double* in) { v[0] = in[0]; v[1] = in[1]; }
};

template <>
struct VarSpec<SE3> {
  static constexpr int32_t kKind = VB_VAR_SE3;
  static constexpr int kData = 7;
  static int tangentDim(const SE3&) { return 6; }
  static uint32_t flags(const SE3&) { return 0u; }
  static double aux(const SE3&) { return 0.0; }
  static void get(const SE3& v, double* o) {
    for (int i = 0; i < 4; i++) o[i] = v.q[i];
    for (int i = 0; i < 3; i++) o[4 + i] = v.t[i];
  }
  static void set(SE3& v, const double* in) {
    for (int i = 0; i < 4; i++) v.q[i] = in[i];
    for (int i = 0; i < 3; i++) v.t[i] = in[4 + i];
  }
};

template <>
struct VarSpec<S2> {
  static constexpr int32_t kKind = VB_VAR_S2;
  static constexpr int kData = 3;
  static int tangentDim(const S2&) { return 2; }
  static uint32_t flags(const S2&) { return 0u; }
  static double aux(const S2& v) { return v.radius; }
  static void get(const S2& v, double* o) { for (int i = 0; i < 3; i++) o[i] = v.vec[i]; }
  static void set(S2& v, const double* in) { for (int i = 0; i < 3; i++) v.vec[i] = in[i]; }
};

template <>
struct VarSpec<CameraModelParam> {
  static constexpr int32_t kKind = VB_VAR_CAM;
  static constexpr int kData = 17;
  static int tangentDim(const CameraModelParam& v) {   // CameraModelParam.h:131-137
    return 15 + (v.estimateReadoutTime ? 1 : 0) + (v.estimateTimeOffset ? 1 : 0);
  }
  static uint32_t flags(const CameraModelParam& v) {
    return (v.estimateReadoutTime ? VB_CAM_EST_READOUT : 0u) | (v.estimateTimeOffset ? VB_CAM_EST_TIMEOFFSET : 0u) |
        (v.hasReadoutTime ? VB_CAM_HAS_READOUT : 0u);
  }
  static double aux(const CameraModelParam&) { return 0.0; }
  static void get(const CameraModelParam& v, double* o) {
    for (int i = 0; i < 15; i++) o[i] = v.params[i];
    o[15] = v.readoutTimeSec;
    o[16] = v.timeOffsetSec;
  }
  static void set(CameraModelParam& v, const double* in) {
    for (int i = 0; i < 15; i++) v.params[i] = in[i];
    v.readoutTimeSec = in[15];
    v.timeOffsetSec = in[16];
  }
};

template <>
struct VarSpec<ImuCalibParam> {
  static constexpr int32_t kKind = VB_VAR_IMUCALIB;
  static constexpr int kData = 32;
  static int tangentDim(const ImuCalibParam& v) { return v.tangentDim; }
  static uint32_t flags(const ImuCalibParam& v) { return v.options; }
  static double aux(const ImuCalibParam&) { return 0.0; }
  static void get(const ImuCalibParam& v, double* o) { for (int i = 0; i < 32; i++) o[i] = v.data[i]; }
  static void set(ImuCalibParam& v, const double* in) { for (int i = 0; i < 32; i++) v.data[i] = in[i]; }
};

// Variable.h:252-330: user-owned value + the index the optimizer assigns
template <typename DType>
struct Variable : public VarBase {
  using DataType = DType;
  Variable() = default;
  explicit Variable(const DType& v) : value(v) {}

  int32_t kind() const override { return VarSpec<DType>::kKind; }
  int getDataDim() const override { return VarSpec<DType>::kData; }
  int getTangentDim() const override { return VarSpec<DType>::tangentDim(value); }
  uint32_t flags() const override { return VarSpec<DType>::flags(value); }
  double aux() const override { return VarSpec<DType>::aux(value); }
  void getData(double* out) const override { VarSpec<DType>::get(value, out); }
  void setData(const double* in) override { VarSpec<DType>::set(value, in); }

  DType value{};
};

}  // namespace small_thing
