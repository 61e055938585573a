// Factor evaluation + assembly kernels: replace
// FactorStore::computeGradHess / computeCost (lib/small_thing/Factor.h:390-417,543-734),
// VariableStore::applyStep (Variable.h:352-370) and Optimizer::addDamping (Optimizer.cpp:136-146).
#include "vb_internal.hpp"
#include "dev_factors_ext.cuh"

namespace vb {

__host__ __device__ inline int errSize(int type) {
  switch (type) {
    case VB_F_VISUAL: case VB_F_VISUAL_RS: return 2;
    case VB_F_INERTIAL: case VB_F_INERTIAL_SEC: case VB_F_INERTIAL_SEC_SPLIT: return 9;
    case VB_F_RW_IMUCALIB: case VB_F_PRIOR_IMUCALIB: return 23;
    case VB_F_RW_CAM: case VB_F_PRIOR_CAM: return 17;
    case VB_F_RW_SE3: case VB_F_PRIOR_SE3: return 6;
    case VB_F_OMEGA_PRIOR: case VB_F_OMEGA_PRIOR_SEC: return 3;
  }
  return 0;
}
__host__ __device__ inline bool hasPrecision(int type) {
  return type == VB_F_INERTIAL || type == VB_F_INERTIAL_SEC || type == VB_F_INERTIAL_SEC_SPLIT;
}
__host__ __device__ inline bool optionalError(int type) { return type == VB_F_VISUAL || type == VB_F_VISUAL_RS; }

// residual (+ Jacobians) of factor k of a low-count store; `type` is launch-uniform
__device__ bool evalSmall(const DevStore& st, const DevVars& V, int64_t k, double* r, double* const* J) {
  const int32_t* v = st.vars + k * st.nvars;
  const double* c = st.consts + k * (int64_t)st.cpf;
#define D(a) (V.data + V.off[v[a]])
  switch (st.type) {
    case VB_F_VISUAL: {
      double JX[6], JT[12], JE[12], JC[30];
      if (!J) return visual_factor<false>(D(0), D(1), D(2), D(3), c, r, nullptr, nullptr, nullptr, nullptr);
      if (!visual_factor<true>(D(0), D(1), D(2), D(3), c, r, JX, JT, JE, JC)) return false;
      const int cd = V.tdim[v[3]];
      if (J[0]) for (int i = 0; i < 6; i++) J[0][i] = JX[i];
      if (J[1]) for (int i = 0; i < 12; i++) J[1][i] = JT[i];
      if (J[2]) for (int i = 0; i < 12; i++) J[2][i] = JE[i];
      if (J[3]) {
        for (int i = 0; i < 2 * cd; i++) J[3][i] = 0;
        for (int j = 0; j < 15; j++) { J[3][j] = JC[j]; J[3][cd + j] = JC[15 + j]; }
      }
      return true;
    }
    case VB_F_VISUAL_RS: {
      double JX[6], JT[12], JE[12], JC[34], JV[6];
      const uint32_t cf = V.flags[v[3]];
      const int status = J ? rs_visual_factor<true>(D(0), D(1), D(2), D(3), cf, D(4), c, st.aux, r, JX, JT, JE, JC, JV)
                           : rs_visual_factor<false>(D(0), D(1), D(2), D(3), cf, D(4), c, st.aux, r, nullptr, nullptr, nullptr, nullptr, nullptr);
      if (status == 2) V.errAcc[0] = 1.0;
      if (status != 0) return false;
      if (!J) return true;
      const int cd = V.tdim[v[3]];
      if (J[0]) for (int i = 0; i < 6; i++) J[0][i] = JX[i];
      if (J[1]) for (int i = 0; i < 12; i++) J[1][i] = JT[i];
      if (J[2]) for (int i = 0; i < 12; i++) J[2][i] = JE[i];
      if (J[3]) {
        for (int e = 0; e < 2; e++) {
          int n = 15;
          for (int j = 0; j < 15; j++) J[3][e * cd + j] = JC[e * 17 + j];
          if (cf & VB_CAM_EST_READOUT) J[3][e * cd + n++] = JC[e * 17 + 15];
          if (cf & VB_CAM_EST_TIMEOFFSET) J[3][e * cd + n++] = JC[e * 17 + 16];
        }
      }
      if (J[4]) for (int i = 0; i < 6; i++) J[4][i] = JV[i];
      return true;
    }
    case VB_F_INERTIAL:
      inertial_factor(D(0), V.flags[v[0]], V.tdim[v[0]], D(1), D(2), D(3), D(4), D(5), st.aux + k * kPreintStride, r, J);
      return true;
    case VB_F_INERTIAL_SEC:
      secondary_inertial_common(D(0), V.flags[v[0]], V.tdim[v[0]], D(1), D(2), D(3), D(4), D(5), D(6), D(7), D(8),
                                st.aux + k * kPreintStride, r, J);
      return true;
    case VB_F_INERTIAL_SEC_SPLIT:
      secondary_inertial(D(0), V.flags[v[0]], V.tdim[v[0]], D(1), D(2), D(3), D(4), D(5), D(6), D(7), D(8), D(9),
                         st.aux + k * kPreintStride, r, J);
      return true;
    case VB_F_RW_IMUCALIB: {
      double d[23];
      for (int i = 0; i < 23; i++) d[i] = 0;
      imu_boxminus(D(1), D(0), V.flags[v[0]], d);
      diag_residual(d, c, V.tdim[v[0]], 23, r, J ? J[0] : nullptr, J ? J[1] : nullptr);
      return true;
    }
    case VB_F_RW_CAM: {
      double d[17];
      for (int i = 0; i < 17; i++) d[i] = 0;
      cam_boxminus(D(1), D(0), V.flags[v[0]], d);
      diag_residual(d, c, V.tdim[v[0]], 17, r, J ? J[0] : nullptr, J ? J[1] : nullptr);
      return true;
    }
    case VB_F_RW_SE3: rw_se3(D(0), D(1), c, r, J); return true;
    case VB_F_PRIOR_IMUCALIB: {
      double d[23];
      for (int i = 0; i < 23; i++) d[i] = 0;
      imu_boxminus(D(0), c + 23, V.flags[v[0]], d);
      diag_residual(d, c, V.tdim[v[0]], 23, r, nullptr, J ? J[0] : nullptr);
      return true;
    }
    case VB_F_PRIOR_CAM: {
      double d[17];
      for (int i = 0; i < 17; i++) d[i] = 0;
      cam_boxminus(D(0), c + 17, V.flags[v[0]], d);
      diag_residual(d, c, V.tdim[v[0]], 17, r, nullptr, J ? J[0] : nullptr);
      return true;
    }
    case VB_F_PRIOR_SE3: prior_se3(D(0), c, r, J); return true;
    case VB_F_OMEGA_PRIOR: omega_prior(D(0), c, r, J); return true;
    case VB_F_OMEGA_PRIOR_SEC: omega_prior_sec(D(0), D(1), c, r, J); return true;
  }
#undef D
  return false;
}

__device__ double squaredError(const DevStore& st, int64_t k, const double* r, int E) {
  double s = 0;
  if (!hasPrecision(st.type)) { for (int i = 0; i < E; i++) s = fma(r[i], r[i], s); return s; }
  const double* L = st.consts + k * (int64_t)st.cpf;
  for (int i = 0; i < E; i++) { double t = 0; for (int j = 0; j < E; j++) t = fma(L[i * E + j], r[j], t); s = fma(r[i], t, s); }
  return s;
}

template <int N>
__device__ void blockReduceAdd(double (&v)[N], double* out) {
  __shared__ double sh[N][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < N; i++) {
    double x = v[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) sh[i][w] = x;
  }
  __syncthreads();
  if (w == 0) {
    const int nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int i = 0; i < N; i++) {
      double x = lane < nw ? sh[i][lane] : 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      if (lane == 0 && x != 0.0) red_add(out + i, x);
    }
  }
}

// ---- generic (low-count) grad/Hess kernel: one warp per factor ------------------------------------
// computeSingleGradHess, Factor.h:543-661.  Lane 0 evaluates residual and Jacobians into shared
// memory; the J~^T J / J~^T r products are then spread over the lanes, one tangent column each.
template <int E, int MAXTOT>
__global__ void __launch_bounds__(128) gradHessSmallKernel(DevStore st, DevVars V, double* grad, double* hess, int updateCache,
                                                            int dontRetry, double* costAcc) {
  __shared__ double Jsh[4][E * MAXTOT];
  __shared__ double rsh[4][E + 2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t idx = blockIdx.x * 4ll + warp;
  const int64_t k = st.active ? (idx < st.nActive ? st.active[idx] : 0) : st.kBegin + idx; // shard: this rank's factors only
  double acc[1] = {0.0};
  if (idx < st.nActive) {
    const int nv = st.nvars;
    const int32_t* v = st.vars + k * nv;
    double* Jb = Jsh[warp];
    double* r = rsh[warp];
    int dims[kMaxVars], joff[kMaxVars];
    int tot = 0;
    for (int i = 0; i < nv; i++) { dims[i] = V.tdim[v[i]]; joff[i] = E * tot; tot += dims[i]; }
    if (lane == 0) {
      double* J[kMaxVars];
      double rl[E];
      for (int i = 0; i < nv; i++) J[i] = st.gOff[k * nv + i] < 0 ? nullptr : Jb + joff[i];
      evalSmall(st, V, k, rl, J);
      double soft, dSoft;
      loss_jet2(st.loss_a, st.loss_k, squaredError(st, k, rl, E), soft, dSoft);
      for (int e = 0; e < E; e++) r[e] = rl[e];
      r[E] = dSoft;
      acc[0] = 0.5 * soft;
    }
    __syncwarp();
    const double dSoft = r[E];
    const double* L = hasPrecision(st.type) ? st.consts + k * (int64_t)st.cpf : nullptr;
    for (int col = lane; col < tot; col += 32) {
      int i = 0, a = col;
      while (a >= dims[i]) { a -= dims[i]; i++; }
      const int64_t gi = st.gOff[k * nv + i];
      if (gi < 0) continue;
      const int di = dims[i];
      const double* Ji = Jb + joff[i];
      double adj[E];
      for (int e = 0; e < E; e++) {
        double s;
        if (L) { s = 0; for (int f = 0; f < E; f++) s = fma(L[e * E + f], Ji[f * di + a], s); }
        else s = Ji[e * di + a];
        adj[e] = s * dSoft;
      }
      double gs = 0;
      for (int e = 0; e < E; e++) gs = fma(r[e], adj[e], gs);
      if (gs != 0.0) red_add(grad + gi + a, gs);
      if (!hess) continue;
      for (int j = 0; j <= i; j++) {
        if (st.gOff[k * nv + j] < 0) continue;
        const int64_t pi = k * st.npairs + i * (i + 1) / 2 + j;
        const int64_t ho = st.hOff[pi];
        if (ho < 0) continue;
        const int sw = st.hStride[pi];
        const int stride = sw & 0x3fffffff;
        const bool tr = (sw >> 30) & 1;
        const int dj = dims[j];
        const double* Jj = Jb + joff[j];
        for (int b = 0; b < dj; b++) {
          double s = 0;
          for (int e = 0; e < E; e++) s = fma(adj[e], Jj[e * dj + b], s);
          if (s != 0.0) red_add(hess + ho + (tr ? (int64_t)b * stride + a : (int64_t)a * stride + b), s);
        }
      }
    }
  }
  blockReduceAdd<1>(acc, costAcc);
}

// ---- visual factors v1 (K1): one CTA per chunk of <= 128 observations sharing rig pose,
// extrinsics and intrinsics.  Thread t evaluates observation t (residual + Jacobians in registers),
// scatters the landmark-side blocks (H_pp, g_p, W_r, W_e, W_c: distinct addresses per landmark)
// with fp64 reductions, and stores its two camera-side Jacobian rows, scaled by sqrt(rho'), plus
// the scaled residual as a 28th column into shared memory:  M = [ sqrt(rho') J_cam | sqrt(rho') r ].
// The camera-side Hessian blocks and gradients of the whole chunk are then ONE contraction
// M^T M (27 x 28, K = 256) on the FP64 tensor pipe, written with one reduction per entry per chunk.
constexpr int kVisChunk = 128;
// LM = false: the landmark-side blocks are left to gradHessVisualLandmarkKernel (two-pass schedule).
template <bool RS, bool LM>
__global__ void __launch_bounds__(128) gradHessVisualChunkKernel(DevStore st, DevVars V, double* grad, double* hess, int updateCache,
                                                                  int dontRetry, double* costAcc) {
  // camera-side columns: pose 0-5, extrinsics 6-11, intrinsics 12-26 (GS) / 12-28 (RS: + readout, time
  // offset), RS only: velocity 29-31; then the residual column
  constexpr int NV = RS ? 5 : 4, NP = NV * (NV + 1) / 2;
  constexpr int CI = RS ? 17 : 15, NC = 12 + CI + (RS ? 3 : 0), LD = RS ? 44 : 36, NCT = RS ? 5 : 4;
  extern __shared__ double M[];          // 256 x LD
  __shared__ double camv[40];            // pose(7) extr(7) intr(17) [vel(3)]
  const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int s0 = st.chunkStart[c], n = st.chunkStart[c + 1] - s0;
  const int f0 = st.order[s0];
  const int32_t* v0 = st.vars + (int64_t)f0 * NV;
  {
    if (tid < 7) camv[tid] = V.data[V.off[v0[1]] + tid];
    else if (tid < 14) camv[tid] = V.data[V.off[v0[2]] + tid - 7];
    else if (tid < 31) camv[tid] = V.data[V.off[v0[3]] + tid - 14];
    else if (RS && tid < 34) camv[tid] = V.data[V.off[v0[4]] + tid - 31];
  }
  const uint32_t camFlags = RS ? V.flags[v0[3]] : 0u;
  const int camDim = V.tdim[v0[3]];
  __syncthreads();
  double acc[1] = {0.0};
  double* row0 = M + (2 * tid) * LD;
  double* row1 = row0 + LD;
  bool live = false;
  if (tid < n) {
    const int f = st.order[s0 + tid];
    if (!(dontRetry && st.cache[f] < 0.0)) {
      double r[2], JX[6], JT[12], JE[12], JC[2 * CI], JV[6];
      const double* X = V.data + V.off[st.vars[(int64_t)f * NV]];
      const double* cst = st.consts + (int64_t)f * st.cpf;
      int status;
      if (RS) status = rs_visual_factor<true>(X, camv, camv + 7, camv + 14, camFlags, camv + 31, cst, st.aux, r, JX, JT, JE, JC, JV);
      else status = visual_factor<true>(X, camv, camv + 7, camv + 14, cst, r, JX, JT, JE, JC) ? 0 : 1;
      if (status == 2) V.errAcc[0] = 1.0;
      if (status != 0) {
        if (updateCache || dontRetry) st.cache[f] = -1.0;
      } else {
        live = true;
        double soft, w;
        loss_jet2(st.loss_a, st.loss_k, r[0] * r[0] + r[1] * r[1], soft, w);
        acc[0] = 0.5 * soft;
        if (updateCache) st.cache[f] = acc[0];
        const double sw = sqrt(w);
#pragma unroll
        for (int j = 0; j < 6; j++) { row0[j] = sw * JT[j]; row1[j] = sw * JT[6 + j]; row0[6 + j] = sw * JE[j]; row1[6 + j] = sw * JE[6 + j]; }
#pragma unroll
        for (int j = 0; j < CI; j++) { row0[12 + j] = sw * JC[j]; row1[12 + j] = sw * JC[CI + j]; }
        if (RS) {
#pragma unroll
          for (int j = 0; j < 3; j++) { row0[12 + CI + j] = sw * JV[j]; row1[12 + CI + j] = sw * JV[3 + j]; }
        }
        row0[NC] = sw * r[0]; row1[NC] = sw * r[1];
#pragma unroll
        for (int j = NC + 1; j < LD; j++) { row0[j] = 0.0; row1[j] = 0.0; }
        // landmark side: weighted J_X
        const int64_t gp = LM ? st.gOff[(int64_t)f * NV] : -1;
        if (LM && gp >= 0) {
          double X0[3], X1[3];
#pragma unroll
          for (int a = 0; a < 3; a++) { X0[a] = w * JX[a]; X1[a] = w * JX[3 + a]; }
#pragma unroll
          for (int a = 0; a < 3; a++) red_add(grad + gp + a, r[0] * X0[a] + r[1] * X1[a]);
          if (hess) {
            const int64_t* ho = st.hOff + (int64_t)f * NP;
            double* Hpp = hess + ho[0];
#pragma unroll
            for (int a = 0; a < 3; a++)
#pragma unroll
              for (int b = 0; b <= a; b++) red_add(Hpp + a * 3 + b, X0[a] * JX[b] + X1[a] * JX[3 + b]);
            if (ho[1] >= 0) {
              double* W = hess + ho[1];
#pragma unroll
              for (int s_ = 0; s_ < 6; s_++)
#pragma unroll
                for (int a = 0; a < 3; a++) red_add(W + s_ * 3 + a, JT[s_] * X0[a] + JT[6 + s_] * X1[a]);
            }
            if (ho[3] >= 0) {
              double* W = hess + ho[3];
#pragma unroll
              for (int s_ = 0; s_ < 6; s_++)
#pragma unroll
                for (int a = 0; a < 3; a++) red_add(W + s_ * 3 + a, JE[s_] * X0[a] + JE[6 + s_] * X1[a]);
            }
            if (ho[6] >= 0) {
              double* W = hess + ho[6];
              // RS: fixed columns 15/16 map to the variable's tangent layout [15, readout?, time offset?]
              const int cRead = 15, cOff = (camFlags & VB_CAM_EST_READOUT) ? 16 : 15;
#pragma unroll
              for (int s_ = 0; s_ < CI; s_++) {
                int row = s_;
                if (RS && s_ == 15) { if (!(camFlags & VB_CAM_EST_READOUT)) continue; row = cRead; }
                if (RS && s_ == 16) { if (!(camFlags & VB_CAM_EST_TIMEOFFSET)) continue; row = cOff; }
#pragma unroll
                for (int a = 0; a < 3; a++) red_add(W + row * 3 + a, JC[s_] * X0[a] + JC[CI + s_] * X1[a]);
              }
            }
            if (RS && ho[10] >= 0) {
              double* W = hess + ho[10];
#pragma unroll
              for (int s_ = 0; s_ < 3; s_++)
#pragma unroll
                for (int a = 0; a < 3; a++) red_add(W + s_ * 3 + a, JV[s_] * X0[a] + JV[3 + s_] * X1[a]);
            }
          }
        }
      }
    }
  }
  if (!live) {
#pragma unroll
    for (int j = 0; j < LD; j++) { row0[j] = 0.0; row1[j] = 0.0; }
  }
  __syncthreads();
  // M^T M on DMMA: warp = 8-row tile of the output, NCT column tiles
  double C[NCT][2];
#pragma unroll
  for (int t = 0; t < NCT; t++) C[t][0] = C[t][1] = 0.0;
  {
    const double* ap = M + (lane & 3) * LD + warp * 8 + (lane >> 2);
    const double* bp = M + (lane & 3) * LD + (lane >> 2);
    for (int k = 0; k < 2 * kVisChunk; k += 4) {
      const double a = ap[k * LD];
#pragma unroll
      for (int t = 0; t < NCT; t++) dmma(C[t][0], C[t][1], a, bp[k * LD + 8 * t]);
    }
  }
  // epilogue: entry (i, j) of [H_cam | g_cam]
  {
    const int i = warp * 8 + (lane >> 2);
    // block of a column index and its offset inside the variable's tangent vector (-1: not estimated)
    auto locate = [&](int col, int& blk, int& a) {
      if (col < 6) { blk = 0; a = col; }
      else if (col < 12) { blk = 1; a = col - 6; }
      else if (col < 12 + CI) {
        blk = 2; a = col - 12;
        if (RS && a == 15) a = (camFlags & VB_CAM_EST_READOUT) ? 15 : -1;
        else if (RS && a == 16) a = (camFlags & VB_CAM_EST_TIMEOFFSET) ? ((camFlags & VB_CAM_EST_READOUT) ? 16 : 15) : -1;
        if (a >= camDim) a = -1;
      } else { blk = 3; a = col - 12 - CI; }
    };
    if (i < NC) {
      int bi, ai;
      locate(i, bi, ai);
      const int64_t gi = st.gOff[(int64_t)f0 * NV + 1 + bi];
      if (gi >= 0 && ai >= 0) {
#pragma unroll
        for (int t = 0; t < NCT; t++)
#pragma unroll
          for (int e = 0; e < 2; e++) {
            const int j = 8 * t + 2 * (lane & 3) + e;
            const double val = C[t][e];
            if (j == NC) { if (val != 0.0) red_add(grad + gi + ai, val); continue; }
            if (!hess || j > i || j > NC) continue;
            int bj, aj;
            locate(j, bj, aj);
            if (aj < 0) continue;
            const int pi = (bi + 1) * (bi + 2) / 2 + (bj + 1);
            const int64_t ho = st.hOff[(int64_t)f0 * NP + pi];
            if (ho < 0) continue;
            const int sw_ = st.hStride[(int64_t)f0 * NP + pi];
            const int stride = sw_ & 0x3fffffff;
            const bool tr = (sw_ >> 30) & 1;
            red_add(hess + ho + (tr ? (int64_t)aj * stride + ai : (int64_t)ai * stride + aj), val);
          }
      }
    }
  }
  blockReduceAdd<1>(acc, costAcc);
}

// ---- visual factors, landmark side (second pass of the two-pass schedule) ---------------------------
// Observations in landmark-major order (sorted by landmark, extrinsics, intrinsics): thread =
// observation, so the lanes of a warp that share a landmark (and, inside it, a calibration window)
// are contiguous.  H_pp, g_p and the landmark x extrinsics / intrinsics blocks are summed inside the
// warp with segmented shuffles and written once per segment; only the landmark x pose (and velocity)
// block, which belongs to a single rig, is written per observation.  ~24 reductions per observation
// instead of ~90 with distinct sectors per lane.
template <int N>
VB_DEV void segReduce(double (&v)[N], int lane, int segEnd) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
    for (int i = 0; i < N; i++) {
      const double t = __shfl_down_sync(0xffffffffu, v[i], d);
      if (lane + d <= segEnd) v[i] += t;
    }
  }
}
VB_DEV int segEndOf(uint32_t heads, int lane) {
  const uint32_t higher = lane == 31 ? 0u : (heads & ~((2u << lane) - 1u));
  return higher ? __ffs(higher) - 2 : 31;
}
template <bool RS>
__global__ void __launch_bounds__(128) gradHessVisualLandmarkKernel(DevStore st, DevVars V, double* grad, double* hess, int dontRetry) {
  constexpr int NV = RS ? 5 : 4, NP = NV * (NV + 1) / 2;
  constexpr int CI = RS ? 17 : 15;
  const int lane = threadIdx.x & 31;
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const bool inRange = i < st.nActive;
  const int f = inRange ? st.orderLm[i] : -1;
  const int32_t* v = st.vars + (int64_t)(inRange ? f : 0) * NV;
  const int vL = inRange ? v[0] : -1 - lane, vE = inRange ? v[2] : -1, vC = inRange ? v[3] : -1;
  // segment heads: landmark segments, and (landmark, extrinsics, intrinsics) segments inside them
  const int pL = __shfl_up_sync(0xffffffffu, vL, 1), pE = __shfl_up_sync(0xffffffffu, vE, 1), pC = __shfl_up_sync(0xffffffffu, vC, 1);
  const bool headL = lane == 0 || pL != vL;
  const bool headC = headL || pE != vE || pC != vC;
  const int endL = segEndOf(__ballot_sync(0xffffffffu, headL), lane);
  const int endC = segEndOf(__ballot_sync(0xffffffffu, headC), lane);
  double r[2] = {0, 0}, JX[6], JT[12], JE[12], JC[2 * CI], JV[6], X0[3] = {0, 0, 0}, X1[3] = {0, 0, 0};
  const uint32_t camFlags = (RS && inRange) ? V.flags[v[3]] : 0u; // heads need the flags even when not live
  bool live = false;
  if (inRange && !(dontRetry && st.cache[f] < 0.0)) {
    const double* cst = st.consts + (int64_t)f * st.cpf;
    const double* X = V.data + V.off[v[0]];
    const double* Tbw = V.data + V.off[v[1]];
    const double* Tcb = V.data + V.off[v[2]];
    const double* cam = V.data + V.off[v[3]];
    int status;
    if (RS) { status = rs_visual_factor<true>(X, Tbw, Tcb, cam, camFlags, V.data + V.off[v[4]], cst, st.aux, r, JX, JT, JE, JC, JV); }
    else status = visual_factor<true>(X, Tbw, Tcb, cam, cst, r, JX, JT, JE, JC) ? 0 : 1;
    live = status == 0;
  }
  if (live) {
    double soft, w;
    loss_jet2(st.loss_a, st.loss_k, r[0] * r[0] + r[1] * r[1], soft, w);
#pragma unroll
    for (int a = 0; a < 3; a++) { X0[a] = w * JX[a]; X1[a] = w * JX[3 + a]; }
  } else {
#pragma unroll
    for (int j = 0; j < 6; j++) JX[j] = 0.0;
#pragma unroll
    for (int j = 0; j < 12; j++) { JT[j] = 0.0; JE[j] = 0.0; }
#pragma unroll
    for (int j = 0; j < 2 * CI; j++) JC[j] = 0.0;
#pragma unroll
    for (int j = 0; j < 6; j++) JV[j] = 0.0;
  }
  const int64_t gp = inRange ? st.gOff[(int64_t)f * NV] : -1;
  const int64_t* ho = st.hOff + (int64_t)(inRange ? f : 0) * NP;
  const bool wr = gp >= 0;
  // gradient + H_pp: one write per landmark segment
  {
    double g[3];
#pragma unroll
    for (int a = 0; a < 3; a++) g[a] = r[0] * X0[a] + r[1] * X1[a];
    segReduce<3>(g, lane, endL);
    if (headL && wr) {
#pragma unroll
      for (int a = 0; a < 3; a++) if (g[a] != 0.0) red_add(grad + gp + a, g[a]);
    }
    if (hess) {
      double h[6];
      int q = 0;
#pragma unroll
      for (int a = 0; a < 3; a++)
#pragma unroll
        for (int b = 0; b <= a; b++) h[q++] = X0[a] * JX[b] + X1[a] * JX[3 + b];
      segReduce<6>(h, lane, endL);
      if (headL && wr) {
        double* Hpp = hess + ho[0];
        q = 0;
#pragma unroll
        for (int a = 0; a < 3; a++)
#pragma unroll
          for (int b = 0; b <= a; b++) { if (h[q] != 0.0) red_add(Hpp + a * 3 + b, h[q]); q++; }
      }
    }
  }
  if (!hess) return;
  // landmark x pose (and velocity): one rig per observation, written per lane
  if (live && wr) {
    if (ho[1] >= 0) {
      double* W = hess + ho[1];
#pragma unroll
      for (int s_ = 0; s_ < 6; s_++)
#pragma unroll
        for (int a = 0; a < 3; a++) red_add(W + s_ * 3 + a, JT[s_] * X0[a] + JT[6 + s_] * X1[a]);
    }
    if (RS && ho[10] >= 0) {
      double* W = hess + ho[10];
#pragma unroll
      for (int s_ = 0; s_ < 3; s_++)
#pragma unroll
        for (int a = 0; a < 3; a++) red_add(W + s_ * 3 + a, JV[s_] * X0[a] + JV[3 + s_] * X1[a]);
    }
  }
  // landmark x extrinsics: one write per (landmark, calibration) segment
  {
    const bool w2 = headC && wr && ho[3] >= 0;
    double* W = hess + (w2 ? ho[3] : 0);
#pragma unroll
    for (int s_ = 0; s_ < 6; s_++) {
      double t[3];
#pragma unroll
      for (int a = 0; a < 3; a++) t[a] = JE[s_] * X0[a] + JE[6 + s_] * X1[a];
      segReduce<3>(t, lane, endC);
      if (w2) {
#pragma unroll
        for (int a = 0; a < 3; a++) if (t[a] != 0.0) red_add(W + s_ * 3 + a, t[a]);
      }
    }
  }
  // landmark x intrinsics (RS: fixed columns 15/16 map to the tangent layout [15, readout?, time offset?])
  {
    const bool w2 = headC && wr && ho[6] >= 0;
    double* W = hess + (w2 ? ho[6] : 0);
    const int cOff = (camFlags & VB_CAM_EST_READOUT) ? 16 : 15;
#pragma unroll
    for (int s_ = 0; s_ < CI; s_++) {
      double t[3];
#pragma unroll
      for (int a = 0; a < 3; a++) t[a] = JC[s_] * X0[a] + JC[CI + s_] * X1[a];
      segReduce<3>(t, lane, endC);
      int row = s_;
      bool on = w2;
      if (RS && s_ == 15) on = on && (camFlags & VB_CAM_EST_READOUT);
      if (RS && s_ == 16) { on = on && (camFlags & VB_CAM_EST_TIMEOFFSET); row = cOff; }
      if (on) {
#pragma unroll
        for (int a = 0; a < 3; a++) if (t[a] != 0.0) red_add(W + row * 3 + a, t[a]);
      }
    }
  }
}

// ---- cost-only kernel: computeSingleCost (Factor.h:390-417) + CostStats ---------------------------
// acc = {cost, numTotal, numInvalid, numPrevInvalid}
template <int E>
__global__ void __launch_bounds__(128) costKernel(DevStore st, DevVars V, int comparable, double* accOut) {
  const int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t k = st.active ? (idx < st.nActive ? st.active[idx] : 0) : st.kBegin + idx; // shard: this rank's factors only
  double acc[4] = {0, 0, 0, 0};
  if (idx < st.nActive) {
    double r[E];
    const bool valid = evalSmall(st, V, k, r, nullptr);
    acc[1] = 1.0;
    bool has = true;
    double val = 0.0;
    if (optionalError(st.type)) {
      const double prev = st.cache[k];
      const bool prevInvalid = prev < 0.0;
      acc[2] = valid ? 0.0 : 1.0;
      acc[3] = prevInvalid ? 1.0 : 0.0;
      if (comparable) {
        if (prevInvalid) has = false;
        else if (!valid) { val = prev; has = false; acc[0] = prev; }
      }
    }
    if (has) {
      if (valid) {
        double soft, d;
        loss_jet2(st.loss_a, st.loss_k, squaredError(st, k, r, E), soft, d);
        acc[0] = 0.5 * soft;
      }
    }
    (void)val;
  }
  blockReduceAdd<4>(acc, accOut);
}

// ---- retraction: VarSpec<T>::applyBoxPlus per variable kind --------------------------------------
// Variable.h:33-37 (Vec), :104-110 (SE3), :195-201 (S2); CameraModelParam.cpp:54-67;
// ImuCalibParam.cpp:55-116,214-221.  ratioAcc = {max (as ordered bits), sum sq, sum, count}
VB_DEV double linf(const double* v, int n) { double m = 0; for (int i = 0; i < n; i++) m = fmax(m, fabs(v[i])); return m; }

__global__ void __launch_bounds__(128) applyStepKernel(DevVars V, const int64_t* blockVar, int64_t nB, const double* step,
                                                        double* ratioAcc) {
  const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  double acc[3] = {0, 0, 0};
  double ratio = 0;
  if (b < nB) {
    const int64_t v = blockVar[b];
    double* x = V.data + V.off[v];
    const double* s = step + V.vecOff[v];
    switch (V.kind[v]) {
      case VB_VAR_VEC3: {
        for (int i = 0; i < 3; i++) x[i] += s[i];
        ratio = linf(s, 3) / (1.0 + linf(x, 3));
        break;
      }
      case VB_VAR_VEC2: {
        for (int i = 0; i < 2; i++) x[i] += s[i];
        ratio = linf(s, 2) / (1.0 + linf(x, 2));
        break;
      }
      case VB_VAR_SE3: {
        double d[6], Ex[7], T0[7], T[7];
        for (int i = 0; i < 6; i++) d[i] = s[i];
        for (int i = 0; i < 7; i++) T0[i] = x[i];
        se3_exp(d, Ex);
        se3_mul(Ex, T0, T);
        for (int i = 0; i < 7; i++) x[i] = T[i];
        ratio = fmax(linf(d + 3, 3), linf(d, 3) / (1.0 + linf(T + 4, 3)));
        break;
      }
      case VB_VAR_S2: {
        const double radius = V.aux[v];
        const double angle = sqrt(s[0] * s[0] + s[1] * s[1]) / radius;
        const double factor = angle > 1e-4 ? (tan(angle) / angle) : (1.0 + angle * angle / 3.0);
        double O[6], xv[3] = {x[0], x[1], x[2]}, nv[3];
        s2_ortho(xv, O);
        for (int i = 0; i < 3; i++) nv[i] = xv[i] + factor * (O[i] * s[0] + O[3 + i] * s[1]);
        const double n = sqrt(dot3(nv, nv));
        for (int i = 0; i < 3; i++) x[i] = nv[i] / n * radius;
        ratio = angle;
        break;
      }
      case VB_VAR_CAM: {
        int n = 15;
        for (int i = 0; i < 15; i++) x[i] += s[i];
        if (V.flags[v] & VB_CAM_EST_READOUT) x[15] += s[n++];
        if (V.flags[v] & VB_CAM_EST_TIMEOFFSET) x[16] += s[n++];
        ratio = linf(s, V.tdim[v]);
        break;
      }
      case VB_VAR_IMUCALIB: {
        double c[32], sl[23];
        for (int i = 0; i < 32; i++) c[i] = x[i];
        for (int i = 0; i < 23; i++) sl[i] = i < V.tdim[v] ? s[i] : 0.0;
        imu_boxplus(c, sl, V.flags[v]);
        for (int i = 0; i < 32; i++) x[i] = c[i];
        ratio = linf(sl, V.tdim[v]);
        break;
      }
    }
    acc[0] = ratio * ratio; acc[1] = ratio; acc[2] = 1.0;
  }
  // max: non-negative doubles order like their bit patterns
  double m = ratio;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0) atomicMax((unsigned long long*)ratioAcc, (unsigned long long)__double_as_longlong(m));
  blockReduceAdd<3>(acc, ratioAcc + 1);
}

// Optimizer::addDamping (Optimizer.cpp:136-146): d <- d (1 + lambda) + lambda on every diagonal scalar
__global__ void dampingKernel(double* hess, const int64_t* diagOff, const int32_t* diagStride, const int32_t* diagSize,
                              int64_t nB, double lambda, const double* addMask) {
  // sharded runs: the diagonals of the shared top are partial sums over ranks, so only rank 0 adds the
  // "+ lambda" term there (the update is affine in d); a subtree's diagonals are complete on its owner,
  // landmark diagonals on theirs (addMask: 1 where this rank adds, null = everywhere)
  const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (b >= nB) return;
  double* D = hess + diagOff[b];
  const int st = diagStride[b], n = diagSize[b];
  if (addMask && addMask[b] < 0.0) return; // a landmark of another rank
  const double add = addMask ? lambda * addMask[b] : lambda;
  for (int a = 0; a < n; a++) { double* p = D + (int64_t)a * st + a; *p = *p * (1.0 + lambda) + add; }
}

// op: 0 a=b ; 1 a*=s ; 2 out[0]+=a.b ; 3 a=0 ; 4 a*=b (elementwise)
__global__ void vecOpKernel(int op, double* a, const double* b, double s, int64_t n, double* out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  double acc[1] = {0.0};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += stride) {
    if (op == 0) a[i] = b[i];
    else if (op == 1) a[i] *= s;
    else if (op == 2) acc[0] = fma(a[i], b[i], acc[0]);
    else if (op == 4) a[i] *= b[i];
    else a[i] = 0.0;
  }
  if (op == 2) blockReduceAdd<1>(acc, out);
}

// ---- launchers ------------------------------------------------------------------------------------
static inline int gridFor(int64_t n, int bs) { return (int)((n + bs - 1) / bs); }

void launchGradHess(const DevStore& st, const DevVars& v, double* grad, double* hess, int updateCache, int dontRetry,
                    double* costAcc, cudaStream_t s) {
  if (st.nActive == 0) return;
#define SMALL(E, TOT) gradHessSmallKernel<E, TOT><<<gridFor(st.nActive, 4), 128, 0, s>>>(st, v, grad, hess, updateCache, dontRetry, costAcc)
  switch (st.type) {
    case VB_F_VISUAL:
    case VB_F_VISUAL_RS: {
      if (!st.order) throw std::runtime_error("visual store without a chunk schedule");
      if (st.nChunks == 0) break; // a shard that owns none of this store's landmarks
      if (st.visImpl) { launchVisualGradHess(st, v, grad, hess, updateCache, dontRetry, costAcc, s); break; }
      const bool rs = st.type == VB_F_VISUAL_RS;
      const size_t smem = (size_t)2 * kVisChunk * (rs ? 44 : 36) * sizeof(double);
      // two passes: camera side per (rig, camera) chunk on DMMA, landmark side in landmark-major order
      // with in-warp segmented sums, when the landmark panels are far larger than L2 (DevStore::twoPass)
      const bool two = st.twoPass && st.orderLm != nullptr;
#define VIS(RS_, LM_)                                                                                                        \
  do {                                                                                                                       \
    VB_CUDA(cudaFuncSetAttribute(gradHessVisualChunkKernel<RS_, LM_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    gradHessVisualChunkKernel<RS_, LM_><<<st.nChunks, 128, smem, s>>>(st, v, grad, hess, updateCache, dontRetry, costAcc);  \
  } while (0)
      if (rs) { if (two) VIS(true, false); else VIS(true, true); }
      else { if (two) VIS(false, false); else VIS(false, true); }
#undef VIS
      if (two) {
        VB_CUDA(cudaGetLastError());
        const int grid = gridFor(st.nActive, 128);
        if (rs) gradHessVisualLandmarkKernel<true><<<grid, 128, 0, s>>>(st, v, grad, hess, dontRetry);
        else gradHessVisualLandmarkKernel<false><<<grid, 128, 0, s>>>(st, v, grad, hess, dontRetry);
      }
      break;
    }
    case VB_F_INERTIAL: SMALL(9, 43); break;
    case VB_F_INERTIAL_SEC: SMALL(9, 55); break;
    case VB_F_INERTIAL_SEC_SPLIT: SMALL(9, 61); break;
    case VB_F_RW_IMUCALIB: SMALL(23, 46); break;
    case VB_F_PRIOR_IMUCALIB: SMALL(23, 23); break;
    case VB_F_RW_CAM: SMALL(17, 34); break;
    case VB_F_PRIOR_CAM: SMALL(17, 17); break;
    case VB_F_RW_SE3: SMALL(6, 12); break;
    case VB_F_PRIOR_SE3: SMALL(6, 6); break;
    case VB_F_OMEGA_PRIOR: SMALL(3, 3); break;
    case VB_F_OMEGA_PRIOR_SEC: SMALL(3, 9); break;
    default: throw std::runtime_error("factor type " + std::to_string(st.type) + " has no CUDA kernel yet");
  }
#undef SMALL
  VB_CUDA(cudaGetLastError());
}

void launchCost(const DevStore& st, const DevVars& v, int comparable, double* acc, cudaStream_t s) {
  if (st.nActive == 0) return;
  if ((st.type == VB_F_VISUAL || st.type == VB_F_VISUAL_RS) && st.visImpl && st.pk.lVar) { launchVisualCost(st, v, comparable, acc, s); return; }
#define COST(E) costKernel<E><<<gridFor(st.nActive, 128), 128, 0, s>>>(st, v, comparable, acc)
  switch (errSize(st.type)) {
    case 2: COST(2); break;
    case 3: COST(3); break;
    case 6: COST(6); break;
    case 9: COST(9); break;
    case 17: COST(17); break;
    case 23: COST(23); break;
    default: throw std::runtime_error("factor type " + std::to_string(st.type) + " has no CUDA kernel yet");
  }
#undef COST
  VB_CUDA(cudaGetLastError());
}

void launchApplyStep(const DevVars& v, const int64_t* blockVar, int64_t nB, const double* step, double* ratioAcc, cudaStream_t s) {
  applyStepKernel<<<gridFor(nB, 128), 128, 0, s>>>(v, blockVar, nB, step, ratioAcc);
  VB_CUDA(cudaGetLastError());
}
void launchDamping(double* hess, const int64_t* diagOff, const int32_t* diagStride, const int32_t* diagSize, int64_t nB,
                   double lambda, const double* addMask, cudaStream_t s) {
  dampingKernel<<<gridFor(nB, 256), 256, 0, s>>>(hess, diagOff, diagStride, diagSize, nB, lambda, addMask);
  VB_CUDA(cudaGetLastError());
}
void launchVecOp(int op, double* a, const double* b, double sc, int64_t n, double* out, cudaStream_t st) {
  if (n <= 0) return;
  const int grid = (int)std::min<int64_t>((n + 255) / 256, 148 * 8);
  vecOpKernel<<<grid, 256, 0, st>>>(op, a, b, sc, n, out);
  VB_CUDA(cudaGetLastError());
}

} // namespace vb
