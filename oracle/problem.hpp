// ORACLE (test infrastructure, CPU only) — see vmath.hpp header note.
//
// Host copy of a vb_graph_desc plus the variable operations of the reference:
// lib/small_thing/Variable.h:25-221 (VarSpec<T>::applyBoxPlus/getData/setData),
// interfaces/ark/camera_model/CameraModelParam.cpp:54-67,
// interfaces/ark/imu_model/ImuCalibParam.cpp:55-116.
#pragma once
#include <cstdint>
#include <vector>
#include <cmath>
#include <algorithm>
#include "../include/viba_b200.h"
#include "vmath.hpp"
#include "imu.hpp"

namespace orc {

struct Store {
  int type, nvars, cpf;
  int64_t count;
  std::vector<int32_t> vars;
  std::vector<double> consts, aux;
  double loss_a, loss_k;
  mutable std::vector<double> cache; // ResultCache (Factor.h:92-94): last valid 0.5*rho, or -1
};

struct Problem {
  int64_t nv = 0, npoints = 0;
  std::vector<int32_t> kind, tdim;
  std::vector<uint32_t> flags;
  std::vector<int64_t> index, off;
  std::vector<double> aux, data, backup;
  std::vector<Store> stores;
  // registration-index -> variable id, sizes
  int64_t nblocks = 0;
  std::vector<int64_t> blockVar;
  std::vector<int32_t> blockSize;

  void load(const vb_graph_desc* g) {
    nv = g->num_vars;
    npoints = g->num_points;
    kind.assign(g->var_kind, g->var_kind + nv);
    tdim.assign(g->var_tdim, g->var_tdim + nv);
    flags.assign(nv, 0);
    if (g->var_flags) flags.assign(g->var_flags, g->var_flags + nv);
    index.assign(g->var_index, g->var_index + nv);
    off.assign(g->var_off, g->var_off + nv);
    aux.assign(nv, 0.0);
    if (g->var_aux) aux.assign(g->var_aux, g->var_aux + nv);
    data.assign(g->var_data, g->var_data + g->var_data_len);
    stores.resize(g->num_stores);
    for (int s = 0; s < g->num_stores; s++) {
      const vb_factor_store& fs = g->stores[s];
      Store& st = stores[s];
      st.type = fs.type; st.nvars = fs.nvars; st.cpf = fs.consts_per_factor; st.count = fs.count;
      st.vars.assign(fs.vars, fs.vars + fs.count * fs.nvars);
      st.consts.assign(fs.consts, fs.consts + fs.count * (int64_t)fs.consts_per_factor);
      if (fs.aux && fs.aux_len > 0) st.aux.assign(fs.aux, fs.aux + fs.aux_len);
      st.loss_a = fs.loss_a; st.loss_k = fs.loss_k;
      st.cache.assign(fs.count, 0.0);
    }
    nblocks = 0;
    for (int64_t v = 0; v < nv; v++) if (index[v] >= 0) nblocks = std::max(nblocks, index[v] + 1);
    blockVar.assign(nblocks, -1);
    blockSize.assign(nblocks, 0);
    for (int64_t v = 0; v < nv; v++) if (index[v] >= 0) { blockVar[index[v]] = v; blockSize[index[v]] = tdim[v]; }
  }

  static int dataDim(int kind) {
    switch (kind) {
      case VB_VAR_VEC3: return 3;
      case VB_VAR_SE3: return 7;
      case VB_VAR_S2: return 3;
      case VB_VAR_CAM: return 17;
      case VB_VAR_IMUCALIB: return 32;
      case VB_VAR_VEC2: return 2;
    }
    return 0;
  }
};

// S2::ortho (Variable.h:168-180): 2x3 row-major
inline void s2Ortho(const double* v, double* res) {
  double t1[3] = {0, 0, 0};
  const double xa = std::fabs(v[0]), ya = std::fabs(v[1]), za = std::fabs(v[2]);
  const int coord = xa < std::min(ya, za) ? 0 : (ya < za ? 1 : 2);
  t1[coord] = 1.0;
  const double vsq = dot3(v, v), vn = std::sqrt(vsq);
  const double c = dot3(t1, v) / vsq;
  double r0[3] = {t1[0] - c * v[0], t1[1] - c * v[1], t1[2] - c * v[2]};
  const double n0 = std::sqrt(dot3(r0, r0));
  for (int i = 0; i < 3; i++) r0[i] /= n0;
  double r1[3];
  cross3(r0, v, r1);
  for (int i = 0; i < 3; i++) { res[i] = r0[i]; res[3 + i] = r1[i] / vn; }
}

inline double linf(const double* v, int n) {
  double m = 0;
  for (int i = 0; i < n; i++) m = std::max(m, std::fabs(v[i]));
  return m;
}

// VarSpec<T>::applyBoxPlus; returns the |step|/|variable| estimate
inline double applyBoxPlus(int kind, uint32_t flags, double aux, double* x, const double* s, int tdim) {
  switch (kind) {
    case VB_VAR_VEC3:
    case VB_VAR_VEC2: { // Variable.h:33-37
      const int n = kind == VB_VAR_VEC3 ? 3 : 2;
      for (int i = 0; i < n; i++) x[i] += s[i];
      return linf(s, n) / (1.0 + linf(x, n));
    }
    case VB_VAR_SE3: { // Variable.h:104-110
      double E[7], T[7];
      se3Exp(s, E);
      se3Mul(E, x, T);
      for (int i = 0; i < 7; i++) x[i] = T[i];
      return std::max(linf(s + 3, 3), linf(s, 3) / (1.0 + linf(x + 4, 3)));
    }
    case VB_VAR_S2: { // Variable.h:195-201
      const double radius = aux;
      const double angle = std::sqrt(s[0] * s[0] + s[1] * s[1]) / radius;
      const double factor = angle > 1e-4 ? (std::tan(angle) / angle) : (1.0 + angle * angle / 3.0);
      double O[6];
      s2Ortho(x, O);
      double v[3];
      for (int i = 0; i < 3; i++) v[i] = x[i] + factor * (O[i] * s[0] + O[3 + i] * s[1]);
      const double n = std::sqrt(dot3(v, v));
      for (int i = 0; i < 3; i++) x[i] = v[i] / n * radius;
      return angle;
    }
    case VB_VAR_CAM: { // CameraModelParam.cpp:54-67
      int n = 15;
      for (int i = 0; i < 15; i++) x[i] += s[i];
      if (flags & VB_CAM_EST_READOUT) x[15] += s[n++];
      if (flags & VB_CAM_EST_TIMEOFFSET) x[16] += s[n++];
      return linf(s, tdim);
    }
    case VB_VAR_IMUCALIB: {
      imuCalibBoxPlus(x, s, flags);
      return linf(s, tdim); // ImuCalibParam.cpp:214-221
    }
  }
  return 0;
}

} // namespace orc
