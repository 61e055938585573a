// Iterative path of the linear solve: lib/small_thing/Optimizer.cpp:232-331 — landmark Schur
// elimination (the kernels of kernels_chol.cu / kernels_solver.cu), then preconditioned conjugate gradients
// (PCG.cpp:15-97) on the reduced camera system, which after the Schur update sits UNFACTORED in the supernode panels:
//   camSymMvKernel     u += S v  (Solver::addMvFrom, Optimizer.cpp:271-283): every stored double of the supernode panels
//                      (lower triangle of the diagonal blocks + below-diagonal panels) read once, used in both directions,
//                      fp64 reductions into u
//   jacobiInitKernel   BlockJacobiPrecond::init (Preconditioner.h:71-82): Cholesky of every param-sized diagonal block
//   jacobiApplyKernel  BlockJacobiPrecond::operator() (:84-107): two triangular solves per block
//   pcg*Kernel         the vector updates and dot products of PCG::solve
// IdentityPrecond (Preconditioner.h:29-47) is a copy.  BlockGaussSeidelPrecond (pseudo-factor: potrf + trsm of every
// supernode without the Schur updates, then the supernodal solves) and LowerPrecSolvePrecond (the full supernodal factor
// of a float-rounded copy) reuse the kernels of kernels_chol.cu on a copy of the camera panels (csrc/api.cu).
#include "vb_internal.hpp"
#include "dev_math.cuh"

namespace vb {

constexpr int kMvRows = 64;

// One CTA per (supernode, slice of kMvRows panel rows): a slice is contiguous in memory (rows x w doubles), staged once
// through shared memory (~50 KB, 4 CTAs per SM so that the copies of three overlap the arithmetic of the fourth) and used
// in both directions.  Row r of the panel belongs to reduced scalar g(r) (inside the diagonal block: the supernode's own
// column r; below: updGlobal); with cl(r) = r inside the diagonal block (lower triangle stored) and w - 1 below,
//   u[g(r)]   += sum_{j <= cl(r)} P[r][j] v[c0 + j]                  (warp per row)
//   u[c0 + j] += P[r][j] v[g(r)]   for j <= cl(r), j != r             (thread per column: the mirrored half)
__global__ void __launch_bounds__(256, 3) camSymMvKernel(DevSolver d, const double* hess, const int2* tasks, const double* v, double* u) {
  extern __shared__ __align__(16) double sm[];
  const int L = tasks[blockIdx.x].x, slice = tasks[blockIdx.x].y;
  const int w = d.lumpWidth[L], wPad = (w + 1) & ~1;
  const int rBeg = slice * kMvRows, n = min(kMvRows, d.lumpRows[L] - rBeg);
  double* xL = sm;                                  // wPad
  double* xr = xL + wPad;                           // kMvRows
  int* gi = reinterpret_cast<int*>(xr + kMvRows);   // kMvRows ints
  double* part = xr + kMvRows + kMvRows / 2;        // 2 wPad
  double* Ps = part + 2 * wPad;                     // kMvRows x w, contiguous
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int c0 = d.camStart[d.lumpFirst[L]];
  const double* src = hess + d.lumpOff[L] + (int64_t)rBeg * w;
  const int tot = n * w;
  if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
    for (int i = 2 * tid; i < tot; i += 512) cp_async16(Ps + i, src + i, 8 * min(2, tot - i));
  } else {
    for (int i = tid; i < tot; i += 256) cp_async8(Ps + i, src + i, true);
  }
  for (int i = tid; i < w; i += 256) xL[i] = v[c0 + i];
  const int64_t q0 = d.updPtr[L];
  for (int i = tid; i < n; i += 256) {
    const int r = rBeg + i;
    const int g = r < w ? c0 + r : d.updGlobal[q0 + r - w];
    gi[i] = g; xr[i] = v[g];
  }
  cp_async_wait_all();
  __syncthreads();
  for (int i = warp; i < n; i += 8) {
    const int r = rBeg + i, cl = r < w ? r : w - 1;
    const double* P0 = Ps + i * w;
    double acc = 0.0;
    for (int j = lane; j <= cl; j += 32) acc = fma(P0[j], xL[j], acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0 && acc != 0.0) red_add(u + gi[i], acc);
  }
  {
    const int half = tid >> 7, c = tid & 127;
    if (c < w) {
      double acc = 0.0;
      for (int i = half; i < n; i += 2) {
        const int r = rBeg + i;
        if (r < w && c >= r) continue;   // diagonal block: only the strictly lower part is mirrored
        acc = fma(Ps[i * w + c], xr[i], acc);
      }
      part[half * wPad + c] = acc;
    }
  }
  __syncthreads();
  for (int c = tid; c < w; c += 256) { const double acc = part[c] + part[wPad + c]; if (acc != 0.0) red_add(u + c0 + c, acc); }
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

// LowerPrecSolvePrecond (Preconditioner.h:160-245): the copy of the camera panels is rounded to float, and so is every
// vector that goes through the preconditioner; the arithmetic on the rounded data stays fp64
__global__ void roundToFloatKernel(double* a, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) a[i] = (double)(float)a[i];
}
// d -> d (1 + eps) + eps on the diagonal of every camera block of the panel copy (Preconditioner.h:200-208)
__global__ void camDiagRaiseKernel(DevSolver d, double* hess, double eps) {
  const int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= d.nC) return;
  const int L = d.camLump[c];
  const int w = d.lumpWidth[L], o = d.camStart[c] - d.camStart[d.lumpFirst[L]], n = d.camStart[c + 1] - d.camStart[c];
  double* D = hess + d.lumpOff[L] + (int64_t)o * w + o;
  for (int a = 0; a < n; a++) D[(int64_t)a * w + a] = D[(int64_t)a * w + a] * (1.0 + eps) + eps;
}

static int gridN(int64_t n) { return (int)std::min<int64_t>((n + 255) / 256, 148 * 8); }

int camSymMvRows() { return kMvRows; }
void launchCamSymMv(const DevSolver& d, const double* hess, const int2* tasks, int nTasks, const double* v, double* u, int maxWidth, cudaStream_t s) {
  if (nTasks <= 0) return;
  const int wPad = (maxWidth + 1) & ~1;
  const size_t smem = ((size_t)3 * wPad + kMvRows + kMvRows / 2 + (size_t)kMvRows * maxWidth) * sizeof(double);
  static size_t cur = 0;
  if (smem > cur) { VB_CUDA(cudaFuncSetAttribute(camSymMvKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); cur = smem; }
  camSymMvKernel<<<nTasks, 256, smem, s>>>(d, hess, tasks, v, u);
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
void launchRoundToFloat(double* a, int64_t n, cudaStream_t s) {
  if (n <= 0) return;
  roundToFloatKernel<<<gridN(n), 256, 0, s>>>(a, n);
  VB_CUDA(cudaGetLastError());
}
void launchCamDiagRaise(const DevSolver& d, double* hess, double eps, cudaStream_t s) {
  camDiagRaiseKernel<<<(unsigned)((d.nC + 127) / 128), 128, 0, s>>>(d, hess, eps);
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
