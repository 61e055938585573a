// Sparse-solver numeric kernels: replace BaSpaCho::Solver::factor / solve as
// consumed at lib/small_thing/Optimizer.cpp:223,228.  Landmark ("sparse elimination") range first:
// per landmark 3x3 Cholesky, panel trsm, Schur update of the camera system; then the supernodal
// Cholesky of the camera system level by level over the elimination tree; then the solves.
#include "vb_internal.hpp"
#include "dev_math.cuh"
#include <cstdio>
#include <cstdlib>

namespace vb {

// row of camera position c inside lump T's panel (device mirror of Symbolic::rowInLump)
VB_DEV int rowInLump(const DevSolver& d, int T, int c) {
  if (d.camLump[c] == T) return d.camStart[c] - d.camStart[d.lumpFirst[T]];
  int64_t lo = d.lumpRowPtr[T], hi = d.lumpRowPtr[T + 1];
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (d.lumpRowCam[mid] < c) lo = mid + 1; else hi = mid;
  }
  return d.lumpRowOff[lo];
}

// ---- K9a: landmark elimination, first half: one warp per landmark factors H_pp (3x3) in place and
// turns the row panel into W L^-T.  The Schur update itself is schurGroupUpdateKernel (kernels_chol.cu).
__global__ void __launch_bounds__(256) schurPrepKernel(DevSolver d, double* hess, int* errFlag) {
  const int64_t ip = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (ip >= d.nLmActive) return;
  const int64_t p = d.lmActive ? d.lmActive[ip] : ip; // shard: the landmarks this rank owns
  double* base = hess + d.lmOff[p];
  const int m = (int)((d.lmOff[p + 1] - d.lmOff[p]) / 3) - 3;
  const double a00 = base[0], a10 = base[3], a11 = base[4], a20 = base[6], a21 = base[7], a22 = base[8];
  const double l00 = sqrt(a00), l10 = a10 / l00, l20 = a20 / l00;
  const double d11 = a11 - l10 * l10;
  const double l11 = sqrt(d11), l21 = (a21 - l20 * l10) / l11;
  const double d22 = a22 - l20 * l20 - l21 * l21;
  const double l22 = sqrt(d22);
  if (!(a00 > 0.0) || !(d11 > 0.0) || !(d22 > 0.0)) { if (lane == 0) atomicExch(errFlag, 1); return; }
  const double i00 = 1.0 / l00, i11 = 1.0 / l11, i22 = 1.0 / l22;
  double* W = base + 9;
  for (int row = lane; row < m; row += 32) { // W <- W L^-T
    double* w = W + row * 3;
    const double x0 = w[0] * i00;
    const double x1 = (w[1] - x0 * l10) * i11;
    const double x2 = (w[2] - x0 * l20 - x1 * l21) * i22;
    w[0] = x0; w[1] = x1; w[2] = x2;
  }
  __syncwarp();
  if (lane == 0) {
    base[0] = l00; base[1] = 0; base[2] = 0; base[3] = l10; base[4] = l11; base[5] = 0; base[6] = l20; base[7] = l21; base[8] = l22;
  }
}

// sharded runs: zero only the landmark panels this rank owns (the others are never touched), one warp
// per landmark
__global__ void __launch_bounds__(256) zeroOwnedPanelsKernel(DevSolver d, double* hess) {
  const int64_t ip = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (ip >= d.nLmActive) return;
  const int64_t p = d.lmActive ? d.lmActive[ip] : ip;
  double* base = hess + d.lmOff[p];
  const int n = (int)(d.lmOff[p + 1] - d.lmOff[p]);
  for (int i = lane; i < n; i += 32) base[i] = 0.0;
}
void launchZeroOwnedPanels(const DevSolver& d, double* hess, cudaStream_t s) {
  if (d.nLmActive == 0) return;
  zeroOwnedPanelsKernel<<<(int)((d.nLmActive * 32 + 255) / 256), 256, 0, s>>>(d, hess);
  VB_CUDA(cudaGetLastError());
}

// ---- K11: solves -------------------------------------------------------------------------------------
// forward over landmarks: y_p = L_pp^-1 g_p ; g_cam -= W'_p y_p
__global__ void __launch_bounds__(256) solveLmFwdKernel(DevSolver d, const double* hess, double* vec) {
  const int64_t ip = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (ip >= d.nLmActive) return;
  const int64_t p = d.lmActive ? d.lmActive[ip] : ip; // shard: the landmarks this rank owns
  const double* base = hess + d.lmOff[p];
  double* g = vec + 3 * p;
  const double y0 = g[0] / base[0];
  const double y1 = (g[1] - base[3] * y0) / base[4];
  const double y2 = (g[2] - base[6] * y0 - base[7] * y1) / base[8];
  __syncwarp();
  if (lane == 0) { g[0] = y0; g[1] = y1; g[2] = y2; }
  const int64_t r0 = d.lmRowPtr[p];
  const int nb = (int)(d.lmRowPtr[p + 1] - r0);
  for (int b = lane; b < nb; b += 32) {
    const int c = d.lmRowCam[r0 + b];
    const int dc = d.camStart[c + 1] - d.camStart[c];
    const double* W = base + (int64_t)d.lmRowOff[r0 + b] * 3;
    double* out = vec + d.camVec0 + d.camStart[c];
    for (int s = 0; s < dc; s++) red_add(out + s, -(W[s * 3] * y0 + W[s * 3 + 1] * y1 + W[s * 3 + 2] * y2));
  }
}
// backward over landmarks: x_p = L_pp^-T (y_p - sum W'^T x_cam)
__global__ void __launch_bounds__(256) solveLmBwdKernel(DevSolver d, const double* hess, double* vec) {
  const int64_t ip = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (ip >= d.nLmActive) return;
  const int64_t p = d.lmActive ? d.lmActive[ip] : ip; // shard: the landmarks this rank owns
  const double* base = hess + d.lmOff[p];
  const int64_t r0 = d.lmRowPtr[p];
  const int nb = (int)(d.lmRowPtr[p + 1] - r0);
  double s0 = 0, s1 = 0, s2 = 0;
  for (int b = lane; b < nb; b += 32) {
    const int c = d.lmRowCam[r0 + b];
    const int dc = d.camStart[c + 1] - d.camStart[c];
    const double* W = base + (int64_t)d.lmRowOff[r0 + b] * 3;
    const double* x = vec + d.camVec0 + d.camStart[c];
    for (int s = 0; s < dc; s++) { s0 = fma(W[s * 3], x[s], s0); s1 = fma(W[s * 3 + 1], x[s], s1); s2 = fma(W[s * 3 + 2], x[s], s2); }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s0 += __shfl_xor_sync(0xffffffffu, s0, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
  if (lane == 0) {
    double* g = vec + 3 * p;
    const double b0 = g[0] - s0, b1 = g[1] - s1, b2 = g[2] - s2;
    const double x2 = b2 / base[8];
    const double x1 = (b1 - base[7] * x2) / base[4];
    const double x0 = (b0 - base[3] * x1 - base[6] * x2) / base[0];
    g[0] = x0; g[1] = x1; g[2] = x2;
  }
}
// ---- FP64 tensor-pipe probe: mma.sync.m8n8k4.f64 issue-bound loop, returns TFLOP/s -------------------
__global__ void dmmaProbeKernel(double* out, int iters) {
  double c0[2] = {0, 0}, c1[2] = {0, 0}, c2[2] = {0, 0}, c3[2] = {0, 0};
  const double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
  for (int i = 0; i < iters; i++) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0[0]), "+d"(c0[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c1[0]), "+d"(c1[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c2[0]), "+d"(c2[1]) : "d"(a), "d"(b));
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c3[0]), "+d"(c3[1]) : "d"(a), "d"(b));
  }
  if (c0[0] + c1[0] + c2[0] + c3[0] + c0[1] + c1[1] + c2[1] + c3[1] == 12345.678) out[0] = 1.0;
}
// the sm_90+ shapes (m16n8k4 / k8 / k16), probed the same way: SHAPE = K; operands per lane A: K/2, B: K/4, C: 4
template <int K>
__global__ void dmmaProbeWideKernel(double* out, int iters) {
  double c[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
  double a[8], b[4];
#pragma unroll
  for (int i = 0; i < 8; i++) a[i] = 1.0 + (threadIdx.x + i) * 1e-9;
#pragma unroll
  for (int i = 0; i < 4; i++) b[i] = 1.0 - (threadIdx.x + i) * 1e-9;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 2; u++) {
      if (K == 4)
        asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+d"(c[u][0]), "+d"(c[u][1]), "+d"(c[u][2]), "+d"(c[u][3]) : "d"(a[0]), "d"(a[1]), "d"(b[0]));
      else if (K == 8)
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+d"(c[u][0]), "+d"(c[u][1]), "+d"(c[u][2]), "+d"(c[u][3])
                     : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
      else
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                     : "+d"(c[u][0]), "+d"(c[u][1]), "+d"(c[u][2]), "+d"(c[u][3])
                     : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                       "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
    }
  }
  double s = 0;
#pragma unroll
  for (int u = 0; u < 2; u++)
#pragma unroll
    for (int i = 0; i < 4; i++) s += c[u][i];
  if (s == 12345.678) out[0] = 1.0;
}
template <typename F>
static double timeProbe(F launch, double flopsPerLaunch, cudaStream_t s) {
  cudaEvent_t e0, e1;
  VB_CUDA(cudaEventCreate(&e0)); VB_CUDA(cudaEventCreate(&e1));
  launch(100);
  VB_CUDA(cudaEventRecord(e0, s));
  launch(20000);
  VB_CUDA(cudaEventRecord(e1, s));
  VB_CUDA(cudaEventSynchronize(e1));
  float ms = 0;
  VB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return flopsPerLaunch / (ms * 1e-3) * 1e-12;
}
// FP64 tensor-pipe peak: the m8n8k4 shape every kernel here uses; VB_VERBOSE also prints the sm_90+ shapes
// m16n8k4 / k8 / k16 measured the same way (perShape[0..3], may be null)
double measureDmmaPeak(cudaStream_t s, double* perShape) {
  double* dummy;
  VB_CUDA(cudaMalloc(&dummy, 8));
  const int blocks = 148 * 4, threads = 256;
  const double warps = (double)blocks * (threads / 32);
  double r[4];
  r[0] = timeProbe([&](int it) { dmmaProbeKernel<<<blocks, threads, 0, s>>>(dummy, it); }, warps * 20000 * 4.0 * (2.0 * 8 * 8 * 4), s);
  r[1] = timeProbe([&](int it) { dmmaProbeWideKernel<4><<<blocks, threads, 0, s>>>(dummy, it); }, warps * 20000 * 2.0 * (2.0 * 16 * 8 * 4), s);
  r[2] = timeProbe([&](int it) { dmmaProbeWideKernel<8><<<blocks, threads, 0, s>>>(dummy, it); }, warps * 20000 * 2.0 * (2.0 * 16 * 8 * 8), s);
  r[3] = timeProbe([&](int it) { dmmaProbeWideKernel<16><<<blocks, threads, 0, s>>>(dummy, it); }, warps * 20000 * 2.0 * (2.0 * 16 * 8 * 16), s);
  VB_CUDA(cudaGetLastError());
  cudaFree(dummy);
  if (perShape) for (int i = 0; i < 4; i++) perShape[i] = r[i];
  if (getenv("VB_VERBOSE"))
    std::fprintf(stderr, "[viba_b200] FP64 mma.sync probes (TFLOP/s): m8n8k4 %.2f, m16n8k4 %.2f, m16n8k8 %.2f, m16n8k16 %.2f\n", r[0], r[1], r[2], r[3]);
  return r[0];
}

// ---- launchers ----------------------------------------------------------------------------------------
void launchSchurPrep(const DevSolver& d, double* hess, int* errFlag, cudaStream_t s) {
  if (d.nLmActive == 0) return;
  schurPrepKernel<<<(int)((d.nLmActive * 32 + 255) / 256), 256, 0, s>>>(d, hess, errFlag);
  VB_CUDA(cudaGetLastError());
}
void launchSolveLandmarksFwd(const DevSolver& d, const double* hess, double* vec, cudaStream_t s) {
  if (d.nLmActive == 0) return;
  solveLmFwdKernel<<<(int)((d.nLmActive * 32 + 255) / 256), 256, 0, s>>>(d, hess, vec);
  VB_CUDA(cudaGetLastError());
}
void launchSolveLandmarksBwd(const DevSolver& d, const double* hess, double* vec, cudaStream_t s) {
  if (d.nLmActive == 0) return;
  solveLmBwdKernel<<<(int)((d.nLmActive * 32 + 255) / 256), 256, 0, s>>>(d, hess, vec);
  VB_CUDA(cudaGetLastError());
}
} // namespace vb
