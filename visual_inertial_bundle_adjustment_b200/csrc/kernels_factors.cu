// Factor evaluation + assembly kernels: replace
// FactorStore::computeGradHess / computeCost (lib/small_thing/Factor.h:390-417,543-734),
// VariableStore::applyStep (Variable.h:352-370) and Optimizer::addDamping (Optimizer.cpp:136-146).
#include "vb_internal.hpp"
#include "dev_factors_small.cuh"

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
    case VB_F_INERTIAL:
      inertial_factor(D(0), V.flags[v[0]], V.tdim[v[0]], D(1), D(2), D(3), D(4), D(5), st.aux + k * kPreintStride, r, J);
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
  const int64_t k = blockIdx.x * 4ll + warp;
  double acc[1] = {0.0};
  if (k < st.count) {
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

// ---- visual factors v0: one thread per observation, fixed dims (3,6,6,15), registers ---------------
__global__ void __launch_bounds__(128) gradHessVisualKernel(DevStore st, DevVars V, double* grad, double* hess, int updateCache,
                                                             int dontRetry, double* costAcc) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  double acc[1] = {0.0};
  if (k < st.count && !(dontRetry && st.cache[k] < 0.0)) {
    const int32_t* v = st.vars + k * 4;
    double r[2], Jall[2][30]; // columns: X 0-2, T 3-8, E 9-14, C 15-29
    {
      double JX[6], JT[12], JE[12], JC[30];
      const bool ok = visual_factor<true>(V.data + V.off[v[0]], V.data + V.off[v[1]], V.data + V.off[v[2]],
                                          V.data + V.off[v[3]], st.consts + k * 6, r, JX, JT, JE, JC);
      if (!ok) {
        if (updateCache || dontRetry) st.cache[k] = -1.0;
        goto done;
      }
#pragma unroll
      for (int e = 0; e < 2; e++) {
#pragma unroll
        for (int j = 0; j < 3; j++) Jall[e][j] = JX[e * 3 + j];
#pragma unroll
        for (int j = 0; j < 6; j++) { Jall[e][3 + j] = JT[e * 6 + j]; Jall[e][9 + j] = JE[e * 6 + j]; }
#pragma unroll
        for (int j = 0; j < 15; j++) Jall[e][15 + j] = JC[e * 15 + j];
      }
    }
    double soft, dSoft;
    loss_jet2(st.loss_a, st.loss_k, r[0] * r[0] + r[1] * r[1], soft, dSoft);
    acc[0] = 0.5 * soft;
    if (updateCache) st.cache[k] = acc[0];
    const int c0[4] = {0, 3, 9, 15}, dd[4] = {3, 6, 6, 15};
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int64_t gi = st.gOff[k * 4 + i];
      if (gi < 0) continue;
#pragma unroll
      for (int a = 0; a < dd[i]; a++) {
        const double a0 = Jall[0][c0[i] + a] * dSoft, a1 = Jall[1][c0[i] + a] * dSoft;
        red_add(grad + gi + a, r[0] * a0 + r[1] * a1);
        if (hess) {
#pragma unroll
          for (int j = 0; j <= i; j++) {
            const int64_t ho = st.hOff[k * 10 + i * (i + 1) / 2 + j];
            if (ho < 0) continue;
            const int sw = st.hStride[k * 10 + i * (i + 1) / 2 + j];
            const int stride = sw & 0x3fffffff;
            const bool tr = (sw >> 30) & 1;
#pragma unroll
            for (int b = 0; b < dd[j]; b++) {
              const double s = a0 * Jall[0][c0[j] + b] + a1 * Jall[1][c0[j] + b];
              red_add(hess + ho + (tr ? (int64_t)b * stride + a : (int64_t)a * stride + b), s);
            }
          }
        }
      }
    }
  }
done:
  blockReduceAdd<1>(acc, costAcc);
}

// ---- visual factors v1 (K1): one CTA per chunk of <= 128 observations sharing rig pose,
// extrinsics and intrinsics.  Thread t evaluates observation t (residual + Jacobians in registers),
// scatters the landmark-side blocks (H_pp, g_p, W_r, W_e, W_c: distinct addresses per landmark)
// with fp64 reductions, and stores its two camera-side Jacobian rows, scaled by sqrt(rho'), plus
// the scaled residual as a 28th column into shared memory:  M = [ sqrt(rho') J_cam | sqrt(rho') r ].
// The camera-side Hessian blocks and gradients of the whole chunk are then ONE contraction
// M^T M (27 x 28, K = 256) on the FP64 tensor pipe, written with one reduction per entry per chunk.
constexpr int kVisChunk = 128, kVisLd = 36;
__global__ void __launch_bounds__(128) gradHessVisualChunkKernel(DevStore st, DevVars V, double* grad, double* hess, int updateCache,
                                                                  int dontRetry, double* costAcc) {
  extern __shared__ double M[];          // 256 x 36
  __shared__ double camv[32];            // pose(7) extr(7) intr(15 used of 17)
  const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int s0 = st.chunkStart[c], n = st.chunkStart[c + 1] - s0;
  const int f0 = st.order[s0];
  {
    const int32_t* v0 = st.vars + (int64_t)f0 * 4;
    if (tid < 7) camv[tid] = V.data[V.off[v0[1]] + tid];
    else if (tid < 14) camv[tid] = V.data[V.off[v0[2]] + tid - 7];
    else if (tid < 29) camv[tid] = V.data[V.off[v0[3]] + tid - 14];
  }
  __syncthreads();
  double acc[1] = {0.0};
  double* row0 = M + (2 * tid) * kVisLd;
  double* row1 = row0 + kVisLd;
  bool live = false;
  if (tid < n) {
    const int f = st.order[s0 + tid];
    if (!(dontRetry && st.cache[f] < 0.0)) {
      double r[2], JX[6], JT[12], JE[12], JC[30];
      const bool ok = visual_factor<true>(V.data + V.off[st.vars[(int64_t)f * 4]], camv, camv + 7, camv + 14,
                                          st.consts + (int64_t)f * 6, r, JX, JT, JE, JC);
      if (!ok) {
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
        for (int j = 0; j < 15; j++) { row0[12 + j] = sw * JC[j]; row1[12 + j] = sw * JC[15 + j]; }
        row0[27] = sw * r[0]; row1[27] = sw * r[1];
#pragma unroll
        for (int j = 28; j < kVisLd; j++) { row0[j] = 0.0; row1[j] = 0.0; }
        // landmark side: weighted J_X
        const int64_t gp = st.gOff[(int64_t)f * 4];
        if (gp >= 0) {
          double X0[3], X1[3];
#pragma unroll
          for (int a = 0; a < 3; a++) { X0[a] = w * JX[a]; X1[a] = w * JX[3 + a]; }
#pragma unroll
          for (int a = 0; a < 3; a++) red_add(grad + gp + a, r[0] * X0[a] + r[1] * X1[a]);
          if (hess) {
            const int64_t* ho = st.hOff + (int64_t)f * 10;
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
#pragma unroll
              for (int s_ = 0; s_ < 15; s_++)
#pragma unroll
                for (int a = 0; a < 3; a++) red_add(W + s_ * 3 + a, JC[s_] * X0[a] + JC[15 + s_] * X1[a]);
            }
          }
        }
      }
    }
  }
  if (!live) {
#pragma unroll
    for (int j = 0; j < kVisLd; j++) { row0[j] = 0.0; row1[j] = 0.0; }
  }
  __syncthreads();
  // M^T M on DMMA: warp = 8-row tile of the output, 4 column tiles
  double C[4][2];
#pragma unroll
  for (int t = 0; t < 4; t++) C[t][0] = C[t][1] = 0.0;
  {
    const double* ap = M + (lane & 3) * kVisLd + warp * 8 + (lane >> 2);
    const double* bp = M + (lane & 3) * kVisLd + (lane >> 2);
    for (int k = 0; k < 2 * kVisChunk; k += 4) {
      const double a = ap[k * kVisLd];
#pragma unroll
      for (int t = 0; t < 4; t++) dmma(C[t][0], C[t][1], a, bp[k * kVisLd + 8 * t]);
    }
  }
  // epilogue: entry (i, j) of [H_cam | g_cam]; rows/cols 0-5 pose, 6-11 extrinsics, 12-26 intrinsics
  {
    const int i = warp * 8 + (lane >> 2);
    if (i < 27) {
      const int bi = i < 6 ? 0 : (i < 12 ? 1 : 2);
      const int ai = i - (bi == 0 ? 0 : (bi == 1 ? 6 : 12));
      const int64_t gi = st.gOff[(int64_t)f0 * 4 + 1 + bi];
      if (gi >= 0) {
#pragma unroll
        for (int t = 0; t < 4; t++)
#pragma unroll
          for (int e = 0; e < 2; e++) {
            const int j = 8 * t + 2 * (lane & 3) + e;
            const double val = C[t][e];
            if (j == 27) { if (val != 0.0) red_add(grad + gi + ai, val); continue; }
            if (!hess || j > i || j >= 27) continue;
            const int bj = j < 6 ? 0 : (j < 12 ? 1 : 2);
            const int aj = j - (bj == 0 ? 0 : (bj == 1 ? 6 : 12));
            const int pi = (bi + 1) * (bi + 2) / 2 + (bj + 1);
            const int64_t ho = st.hOff[(int64_t)f0 * 10 + pi];
            if (ho < 0) continue;
            const int sw_ = st.hStride[(int64_t)f0 * 10 + pi];
            const int stride = sw_ & 0x3fffffff;
            const bool tr = (sw_ >> 30) & 1;
            red_add(hess + ho + (tr ? (int64_t)aj * stride + ai : (int64_t)ai * stride + aj), val);
          }
      }
    }
  }
  blockReduceAdd<1>(acc, costAcc);
}

// ---- cost-only kernel: computeSingleCost (Factor.h:390-417) + CostStats ---------------------------
// acc = {cost, numTotal, numInvalid, numPrevInvalid}
template <int E>
__global__ void __launch_bounds__(128) costKernel(DevStore st, DevVars V, int comparable, double* accOut) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  double acc[4] = {0, 0, 0, 0};
  if (k < st.count) {
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
                              int64_t nB, double lambda) {
  const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (b >= nB) return;
  double* D = hess + diagOff[b];
  const int st = diagStride[b], n = diagSize[b];
  for (int a = 0; a < n; a++) { double* p = D + (int64_t)a * st + a; *p = *p * (1.0 + lambda) + lambda; }
}

// op: 0 a=b ; 1 a*=s ; 2 out[0]+=a.b ; 3 a=0
__global__ void vecOpKernel(int op, double* a, const double* b, double s, int64_t n, double* out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  double acc[1] = {0.0};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += stride) {
    if (op == 0) a[i] = b[i];
    else if (op == 1) a[i] *= s;
    else if (op == 2) acc[0] = fma(a[i], b[i], acc[0]);
    else a[i] = 0.0;
  }
  if (op == 2) blockReduceAdd<1>(acc, out);
}

// ---- launchers ------------------------------------------------------------------------------------
static inline int gridFor(int64_t n, int bs) { return (int)((n + bs - 1) / bs); }

void launchGradHess(const DevStore& st, const DevVars& v, double* grad, double* hess, int updateCache, int dontRetry,
                    double* costAcc, cudaStream_t s) {
  if (st.count == 0) return;
#define SMALL(E, TOT) gradHessSmallKernel<E, TOT><<<gridFor(st.count, 4), 128, 0, s>>>(st, v, grad, hess, updateCache, dontRetry, costAcc)
  switch (st.type) {
    case VB_F_VISUAL:
      if (st.order) {
        const size_t smem = (size_t)2 * kVisChunk * kVisLd * sizeof(double);
        static bool attr = false;
        if (!attr) { VB_CUDA(cudaFuncSetAttribute(gradHessVisualChunkKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = true; }
        gradHessVisualChunkKernel<<<st.nChunks, 128, smem, s>>>(st, v, grad, hess, updateCache, dontRetry, costAcc);
      } else {
        gradHessVisualKernel<<<gridFor(st.count, 128), 128, 0, s>>>(st, v, grad, hess, updateCache, dontRetry, costAcc);
      }
      break;
    case VB_F_INERTIAL: SMALL(9, 43); break;
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
  if (st.count == 0) return;
#define COST(E) costKernel<E><<<gridFor(st.count, 128), 128, 0, s>>>(st, v, comparable, acc)
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
                   double lambda, cudaStream_t s) {
  dampingKernel<<<gridFor(nB, 256), 256, 0, s>>>(hess, diagOff, diagStride, diagSize, nB, lambda);
  VB_CUDA(cudaGetLastError());
}
void launchVecOp(int op, double* a, const double* b, double sc, int64_t n, double* out, cudaStream_t st) {
  if (n <= 0) return;
  const int grid = (int)std::min<int64_t>((n + 255) / 256, 148 * 8);
  vecOpKernel<<<grid, 256, 0, st>>>(op, a, b, sc, n, out);
  VB_CUDA(cudaGetLastError());
}

} // namespace vb
