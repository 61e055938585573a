// Iterative path of the linear solve: lib/small_thing/Optimizer.cpp:232-331 — landmark Schur
// elimination (the kernels of kernels_chol.cu / kernels_solver.cu), then preconditioned conjugate gradients
// (PCG.cpp:15-97) on the reduced camera system, which after the Schur update sits UNFACTORED in the supernode panels:
//   camSymMvKernel     u += S v  (Solver::addMvFrom, Optimizer.cpp:271-283): per supernode the symmetric diagonal block and
//                      the below-diagonal panel, used in both directions, fp64 reductions into u
//   jacobiInitKernel   BlockJacobiPrecond::init (Preconditioner.h:71-82): Cholesky of every param-sized diagonal block
//   jacobiApplyKernel  BlockJacobiPrecond::operator() (:84-107): two triangular solves per block
//   pcg*Kernel         the vector updates and dot products of PCG::solve
// IdentityPrecond (Preconditioner.h:29-47) is a copy.  BlockGaussSeidelPrecond / LowerPrecSolvePrecond are BaSpaCho's
// pseudoFactorFrom / float factorFrom: source absent from the reference tree, not built.
#include "vb_internal.hpp"
#include "dev_math.cuh"

namespace vb {

constexpr int kMvSlice = 128;

__global__ void __launch_bounds__(256) camSymMvKernel(DevSolver d, const double* hess, const double* v, double* u) {
  extern __shared__ double sm[];
  const int L = blockIdx.y;
  const int w = d.lumpWidth[L], r = d.lumpRows[L] - w;
  const int slice = blockIdx.x;
  const int nSlices = max(1, (r + kMvSlice - 1) / kMvSlice);
  if (slice >= nSlices) return;
  double* xL = sm;                 // w
  double* xb = xL + w;             // kMvSlice
  double* part = xb + kMvSlice;    // 2 w
  double* Ps = part + 2 * w;       // max(w, kMvSlice) x w
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int c0 = d.camStart[d.lumpFirst[L]];
  const double* panel = hess + d.lumpOff[L];
  for (int i = tid; i < w; i += 256) xL[i] = v[c0 + i];
  if (slice == 0) {   // symmetric diagonal block (lower triangle stored)
    loadTileAsync(Ps, w, w, w, panel, w, w, w);
    cp_async_wait_all();
    __syncthreads();
    for (int i = warp; i < w; i += 8) {
      double s = 0.0;
      for (int j = lane; j < w; j += 32) s = fma(j <= i ? Ps[i * w + j] : Ps[j * w + i], xL[j], s);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0 && s != 0.0) red_add(u + c0 + i, s);
    }
    __syncthreads();
  } else {
    __syncthreads();
  }
  const int rBeg = slice * kMvSlice, rEnd = min(r, rBeg + kMvSlice);
  if (rEnd <= rBeg) return;
  const double* below = panel + (int64_t)w * w;
  const int64_t q0 = d.updPtr[L];
  loadTileAsync(Ps, w, rEnd - rBeg, w, below + (int64_t)rBeg * w, w, rEnd - rBeg, w);
  for (int i = tid; i < rEnd - rBeg; i += 256) xb[i] = v[d.updGlobal[q0 + rBeg + i]];
  cp_async_wait_all();
  __syncthreads();
  // u_below += P v_L : warp per row
  for (int row = warp; row < rEnd - rBeg; row += 8) {
    const double* P0 = Ps + row * w;
    double s = 0.0;
    for (int c = lane; c < w; c += 32) s = fma(P0[c], xL[c], s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0 && s != 0.0) red_add(u + d.updGlobal[q0 + rBeg + row], s);
  }
  // u_L += P^T v_below : thread per column, rows split over the two halves of the CTA
  {
    const int half = tid >> 7, c = tid & 127;
    if (c < w) {
      double s = 0.0;
      for (int row = half; row < rEnd - rBeg; row += 2) s = fma(Ps[row * w + c], xb[row], s);
      part[half * w + c] = s;
    }
  }
  __syncthreads();
  for (int c = tid; c < w; c += 256) { const double s = part[c] + part[w + c]; if (s != 0.0) red_add(u + c0 + c, s); }
}

// per camera-side param block: Cholesky factor of its diagonal block of the reduced system (<= 23 x 23), lower, row-major
__global__ void jacobiInitKernel(DevSolver d, const double* hess, const int64_t* jacOff, double* jac, int* errFlag) {
  const int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= d.nC) return;
  const int L = d.camLump[c];
  const int w = d.lumpWidth[L], o = d.camStart[c] - d.camStart[d.lumpFirst[L]], n = d.camStart[c + 1] - d.camStart[c];
  const double* D = hess + d.lumpOff[L] + (int64_t)o * w + o;
  double* Lc = jac + jacOff[c];
  for (int a = 0; a < n; a++) {
    for (int b = 0; b < a; b++) {
      double t = D[(int64_t)a * w + b];
      for (int k = 0; k < b; k++) t = fma(-Lc[a * n + k], Lc[b * n + k], t);
      Lc[a * n + b] = t / Lc[b * n + b];
    }
    double t = D[(int64_t)a * w + a];
    for (int k = 0; k < a; k++) t = fma(-Lc[a * n + k], Lc[a * n + k], t);
    if (!(t > 0.0)) { atomicExch(errFlag, 2); t = 1.0; }
    Lc[a * n + a] = sqrt(t);
  }
}
__global__ void jacobiApplyKernel(DevSolver d, const int64_t* jacOff, const double* jac, const double* in, double* out) {
  const int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= d.nC) return;
  const int s0 = d.camStart[c], n = d.camStart[c + 1] - s0;
  const double* Lc = jac + jacOff[c];
  double x[24];
  for (int a = 0; a < n; a++) {
    double t = in[s0 + a];
    for (int k = 0; k < a; k++) t = fma(-Lc[a * n + k], x[k], t);
    x[a] = t / Lc[a * n + a];
  }
  for (int a = n - 1; a >= 0; a--) {
    double t = x[a];
    for (int k = a + 1; k < n; k++) t = fma(-Lc[k * n + a], x[k], t);
    x[a] = t / Lc[a * n + a];
  }
  for (int a = 0; a < n; a++) out[s0 + a] = x[a];
}

template <int N>
__device__ void pcgBlockReduce(double (&v)[N], double* out) {
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
#pragma unroll
    for (int i = 0; i < N; i++) {
      double x = lane < (blockDim.x >> 5) ? sh[i][lane] : 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      if (lane == 0) red_add(out + i, x);
    }
  }
}
// out[0] += a.b ; out[1] += c.d
__global__ void pcgDot2Kernel(const double* a, const double* b, const double* c, const double* dd, int64_t n, double* out) {
  double acc[2] = {0.0, 0.0};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    acc[0] = fma(a[i], b[i], acc[0]);
    acc[1] = fma(c[i], dd[i], acc[1]);
  }
  pcgBlockReduce<2>(acc, out);
}
// x += alpha p ; r -= alpha Ap   (alpha = num[0] / den[0] read on the device: no host round trip between the two dots)
__global__ void pcgStepKernel(double* x, double* r, const double* p, const double* Ap, const double* num, const double* den, int64_t n) {
  const double alpha = num[0] / den[0];
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    x[i] = fma(alpha, p[i], x[i]);
    r[i] = fma(-alpha, Ap[i], r[i]);
  }
}
// p = z + beta p   (beta = num[0] / den[0])
__global__ void pcgDirKernel(double* p, const double* z, const double* num, const double* den, int64_t n) {
  const double beta = num[0] / den[0];
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = fma(beta, p[i], z[i]);
}

static int gridN(int64_t n) { return (int)std::min<int64_t>((n + 255) / 256, 148 * 8); }

void launchCamSymMv(const DevSolver& d, const double* hess, const double* v, double* u, int nLumps, int maxBelow, int maxWidth, cudaStream_t s) {
  if (nLumps <= 0) return;
  const int slices = std::max(1, (maxBelow + kMvSlice - 1) / kMvSlice);
  const size_t smem = ((size_t)3 * maxWidth + kMvSlice + (size_t)std::max(maxWidth, kMvSlice) * maxWidth) * sizeof(double);
  static size_t cur = 0;
  if (smem > cur) { VB_CUDA(cudaFuncSetAttribute(camSymMvKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); cur = smem; }
  camSymMvKernel<<<dim3(slices, nLumps), 256, smem, s>>>(d, hess, v, u);
  VB_CUDA(cudaGetLastError());
}
void launchJacobiInit(const DevSolver& d, const double* hess, const int64_t* jacOff, double* jac, int* errFlag, cudaStream_t s) {
  jacobiInitKernel<<<(unsigned)((d.nC + 127) / 128), 128, 0, s>>>(d, hess, jacOff, jac, errFlag);
  VB_CUDA(cudaGetLastError());
}
void launchJacobiApply(const DevSolver& d, const int64_t* jacOff, const double* jac, const double* in, double* out, cudaStream_t s) {
  jacobiApplyKernel<<<(unsigned)((d.nC + 127) / 128), 128, 0, s>>>(d, jacOff, jac, in, out);
  VB_CUDA(cudaGetLastError());
}
void launchPcgDot2(const double* a, const double* b, const double* c, const double* dd, int64_t n, double* out, cudaStream_t s) {
  pcgDot2Kernel<<<gridN(n), 256, 0, s>>>(a, b, c, dd, n, out);
  VB_CUDA(cudaGetLastError());
}
void launchPcgStep(double* x, double* r, const double* p, const double* Ap, const double* num, const double* den, int64_t n, cudaStream_t s) {
  pcgStepKernel<<<gridN(n), 256, 0, s>>>(x, r, p, Ap, num, den, n);
  VB_CUDA(cudaGetLastError());
}
void launchPcgDir(double* p, const double* z, const double* num, const double* den, int64_t n, cudaStream_t s) {
  pcgDirKernel<<<gridN(n), 256, 0, s>>>(p, z, num, den, n);
  VB_CUDA(cudaGetLastError());
}

}  // namespace vb
