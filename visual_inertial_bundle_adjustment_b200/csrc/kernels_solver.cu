// Sparse-solver numeric kernels: replace BaSpaCho::Solver::factor / solve as
// consumed at lib/small_thing/Optimizer.cpp:223,228.  Landmark ("sparse elimination") range first:
// per landmark 3x3 Cholesky, panel trsm, Schur update of the camera system; then the supernodal
// Cholesky of the camera system level by level over the elimination tree; then the solves.
#include "vb_internal.hpp"
#include "dev_math.cuh"

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

// ---- K9: landmark elimination, one warp per landmark ------------------------------------------------
__global__ void __launch_bounds__(256) schurKernel(DevSolver d, double* hess, int* errFlag) {
  const int64_t p = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (p >= d.nP) return;
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
  const int64_t r0 = d.lmRowPtr[p];
  const int nb = (int)(d.lmRowPtr[p + 1] - r0);
  const int npairs = nb * (nb + 1) / 2;
  for (int pr = lane; pr < npairs; pr += 32) {
    int ia = (int)((sqrt(8.0 * pr + 1.0) - 1.0) * 0.5);
    while ((ia + 1) * (ia + 2) / 2 <= pr) ia++;
    while (ia * (ia + 1) / 2 > pr) ia--;
    const int ib = pr - ia * (ia + 1) / 2;
    const int ca = d.lmRowCam[r0 + ia], cb = d.lmRowCam[r0 + ib];
    const int da = d.camStart[ca + 1] - d.camStart[ca], db = d.camStart[cb + 1] - d.camStart[cb];
    const int T = d.camLump[cb];
    const int wT = d.lumpWidth[T];
    const int col = d.camStart[cb] - d.camStart[d.lumpFirst[T]];
    const int row = rowInLump(d, T, ca);
    double* tgt = hess + d.lumpOff[T] + (int64_t)row * wT + col;
    const double* Wa = W + (d.lmRowOff[r0 + ia] - 3) * 3;
    const double* Wb = W + (d.lmRowOff[r0 + ib] - 3) * 3;
    for (int s = 0; s < da; s++) {
      const double x0 = Wa[s * 3], x1 = Wa[s * 3 + 1], x2 = Wa[s * 3 + 2];
      const int tmax = (ia == ib) ? s + 1 : db;
      for (int t = 0; t < tmax; t++)
        red_add(tgt + (int64_t)s * wT + t, -(x0 * Wb[t * 3] + x1 * Wb[t * 3 + 1] + x2 * Wb[t * 3 + 2]));
    }
  }
}

// ---- K10a: diagonal potrf (in shared memory) + panel trsm ------------------------------------------
// grid = (row tiles, lumps of the level); every CTA factors the diagonal block redundantly (no
// inter-CTA dependency), CTA x==0 publishes L into diagL, each CTA solves its 64 panel rows.
constexpr int kTrsmRows = 64;
__global__ void __launch_bounds__(128) lumpFactorKernel(DevSolver d, double* hess, const int32_t* lumps, int* errFlag) {
  extern __shared__ double sm[];
  const int L = lumps[blockIdx.y];
  const int w = d.lumpWidth[L], r = d.lumpRows[L] - w;
  const int tile = blockIdx.x;
  if (tile > 0 && tile * kTrsmRows >= r) return;
  const int ld = w + 1;
  double* A = sm;                 // w x ld
  double* X = sm + (size_t)w * ld; // kTrsmRows x ld
  double* panel = hess + d.lumpOff[L];
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < w * w; i += nt) A[(i / w) * ld + (i % w)] = panel[i];
  const int row0 = tile * kTrsmRows;
  const int nrows = min(kTrsmRows, r - row0);
  const double* below = panel + (int64_t)(w + row0) * w;
  for (int i = tid; i < nrows * w; i += nt) X[(i / w) * ld + (i % w)] = below[i];
  __syncthreads();
  for (int j = 0; j < w; j++) {
    const double djj = A[j * ld + j];
    if (!(djj > 0.0)) { if (tid == 0) atomicExch(errFlag, 2); return; } // uniform: every thread reads the same value
    const double ljj = sqrt(djj), inv = 1.0 / ljj;
    __syncthreads();
    for (int i = j + tid; i < w; i += nt) A[i * ld + j] = (i == j) ? ljj : A[i * ld + j] * inv;
    __syncthreads();
    // trailing update of the lower triangle: row i, columns j+1..i
    for (int i = j + 1 + tid; i < w; i += nt) {
      const double lij = A[i * ld + j];
      for (int k = j + 1; k <= i; k++) A[i * ld + k] -= lij * A[k * ld + j];
    }
    __syncthreads();
  }
  if (tile == 0) {
    double* out = d.diagL + d.diagLOff[L];
    for (int i = tid; i < w * w; i += nt) { const int a = i / w, b = i % w; out[i] = b <= a ? A[a * ld + b] : 0.0; }
  }
  if (tid < nrows) { // x L^T = b, one thread per panel row
    double* x = X + tid * ld;
    for (int c = 0; c < w; c++) {
      double s = x[c];
      const double* Lc = A + c * ld;
      for (int k = 0; k < c; k++) s = fma(-x[k], Lc[k], s);
      x[c] = s / Lc[c];
    }
  }
  __syncthreads();
  double* belowW = panel + (int64_t)(w + row0) * w;
  for (int i = tid; i < nrows * w; i += nt) belowW[i] = X[(i / w) * ld + (i % w)];
}

// ---- K10b: supernode update C = B B^T scattered into the ancestors (v0: SIMT 64x64 tiles) ----------
constexpr int kUpdTile = 64;
__global__ void __launch_bounds__(256) lumpUpdateKernel(DevSolver d, double* hess, const LevelTask* tasks, int useAtomics) {
  extern __shared__ double sm[];
  const LevelTask t = tasks[blockIdx.x];
  const int L = t.lump;
  const int w = d.lumpWidth[L], r = d.lumpRows[L] - w;
  const int ld = w + 1;
  double* As = sm;                            // rows of tile ti
  double* Bs = sm + (size_t)kUpdTile * ld;    // rows of tile tj
  const double* below = hess + d.lumpOff[L] + (int64_t)w * w;
  const int i0 = t.ti * kUpdTile, j0 = t.tj * kUpdTile;
  const int ni = min(kUpdTile, r - i0), nj = min(kUpdTile, r - j0);
  const int tid = threadIdx.x;
  for (int i = tid; i < kUpdTile * w; i += 256) {
    const int a = i / w, b = i % w;
    As[a * ld + b] = a < ni ? below[(int64_t)(i0 + a) * w + b] : 0.0;
    Bs[a * ld + b] = a < nj ? below[(int64_t)(j0 + a) * w + b] : 0.0;
  }
  __syncthreads();
  const int ty = tid >> 4, tx = tid & 15; // 16x16 threads, 4x4 micro-tiles
  double acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; a++)
#pragma unroll
    for (int b = 0; b < 4; b++) acc[a][b] = 0.0;
  for (int k = 0; k < w; k++) {
    double av[4], bv[4];
#pragma unroll
    for (int a = 0; a < 4; a++) { av[a] = As[(ty + 16 * a) * ld + k]; bv[a] = Bs[(tx + 16 * a) * ld + k]; }
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
      for (int b = 0; b < 4; b++) acc[a][b] = fma(av[a], bv[b], acc[a][b]);
  }
  const int64_t q0 = d.updPtr[L];
  const int32_t* rel = d.rel + d.relPtr[L];
#pragma unroll
  for (int b = 0; b < 4; b++) {
    const int qj = j0 + tx + 16 * b;
    if (qj >= r) continue;
    const int64_t colOff = d.updColOff[q0 + qj];
    const int stride = d.updStride[q0 + qj];
    const int32_t* relT = rel + (int64_t)d.updTgt[q0 + qj] * r;
#pragma unroll
    for (int a = 0; a < 4; a++) {
      const int qi = i0 + ty + 16 * a;
      if (qi >= r || qi < qj) continue;
      double* p = hess + colOff + (int64_t)relT[qi] * stride;
      if (useAtomics) red_add(p, -acc[a][b]); else *p -= acc[a][b];
    }
  }
}

// ---- K11: solves -------------------------------------------------------------------------------------
// forward over landmarks: y_p = L_pp^-1 g_p ; g_cam -= W'_p y_p
__global__ void __launch_bounds__(256) solveLmFwdKernel(DevSolver d, const double* hess, double* vec) {
  const int64_t p = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (p >= d.nP) return;
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
  const int64_t p = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (p >= d.nP) return;
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
// forward over the lumps of one level: x_L = L^-1 b_L ; b_below -= P x_L.  One CTA per lump.
__global__ void __launch_bounds__(128) solveLumpFwdKernel(DevSolver d, const double* hess, double* vec, const int32_t* lumps, int useAtomics) {
  extern __shared__ double sm[];
  const int L = lumps[blockIdx.x];
  const int w = d.lumpWidth[L], r = d.lumpRows[L] - w;
  double* x = sm;
  double* vL = vec + d.camVec0 + d.camStart[d.lumpFirst[L]];
  const double* Ld = d.diagL + d.diagLOff[L];
  const int tid = threadIdx.x;
  for (int i = tid; i < w; i += 128) x[i] = vL[i];
  __syncthreads();
  for (int c = 0; c < w; c++) {
    if (tid == 0) x[c] /= Ld[c * w + c];
    __syncthreads();
    const double xc = x[c];
    for (int k = c + 1 + tid; k < w; k += 128) x[k] = fma(-Ld[k * w + c], xc, x[k]);
    __syncthreads();
  }
  for (int i = tid; i < w; i += 128) vL[i] = x[i];
  const double* below = hess + d.lumpOff[L] + (int64_t)w * w;
  const int64_t q0 = d.updPtr[L];
  for (int row = tid; row < r; row += 128) {
    const double* P = below + (int64_t)row * w;
    double s = 0;
    for (int c = 0; c < w; c++) s = fma(P[c], x[c], s);
    double* out = vec + d.camVec0 + d.updGlobal[q0 + row];
    if (useAtomics) red_add(out, -s); else *out -= s;
  }
}
// backward: x_L = L^-T (b_L - P^T x_below)
__global__ void __launch_bounds__(128) solveLumpBwdKernel(DevSolver d, const double* hess, double* vec, const int32_t* lumps) {
  extern __shared__ double sm[];
  const int L = lumps[blockIdx.x];
  const int w = d.lumpWidth[L], r = d.lumpRows[L] - w;
  double* x = sm;       // w
  double* xb = sm + w;  // r
  double* vL = vec + d.camVec0 + d.camStart[d.lumpFirst[L]];
  const double* Ld = d.diagL + d.diagLOff[L];
  const double* below = hess + d.lumpOff[L] + (int64_t)w * w;
  const int64_t q0 = d.updPtr[L];
  const int tid = threadIdx.x;
  for (int row = tid; row < r; row += 128) xb[row] = vec[d.camVec0 + d.updGlobal[q0 + row]];
  __syncthreads();
  for (int c = tid; c < w; c += 128) {
    double s = vL[c];
    for (int row = 0; row < r; row++) s = fma(-below[(int64_t)row * w + c], xb[row], s);
    x[c] = s;
  }
  __syncthreads();
  for (int c = w - 1; c >= 0; c--) {
    if (tid == 0) x[c] /= Ld[c * w + c];
    __syncthreads();
    const double xc = x[c];
    for (int k = tid; k < c; k += 128) x[k] = fma(-Ld[c * w + k], xc, x[k]);
    __syncthreads();
  }
  for (int i = tid; i < w; i += 128) vL[i] = x[i];
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
double measureDmmaPeak(cudaStream_t s) {
  double* dummy;
  VB_CUDA(cudaMalloc(&dummy, 8));
  const int iters = 20000, blocks = 148 * 4, threads = 256;
  dmmaProbeKernel<<<blocks, threads, 0, s>>>(dummy, 100);
  cudaEvent_t e0, e1;
  VB_CUDA(cudaEventCreate(&e0)); VB_CUDA(cudaEventCreate(&e1));
  VB_CUDA(cudaEventRecord(e0, s));
  dmmaProbeKernel<<<blocks, threads, 0, s>>>(dummy, iters);
  VB_CUDA(cudaEventRecord(e1, s));
  VB_CUDA(cudaEventSynchronize(e1));
  float ms = 0;
  VB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(dummy);
  const double flops = (double)blocks * (threads / 32) * iters * 4.0 * (2.0 * 8 * 8 * 4);
  return flops / (ms * 1e-3) * 1e-12;
}

// ---- launchers ----------------------------------------------------------------------------------------
void launchSchur(const DevSolver& d, double* hess, int* errFlag, cudaStream_t s) {
  if (d.nP == 0) return;
  schurKernel<<<(int)((d.nP * 32 + 255) / 256), 256, 0, s>>>(d, hess, errFlag);
  VB_CUDA(cudaGetLastError());
}
static void ensureSmem(const void* fn, size_t bytes) {
  if (bytes > 48 * 1024) VB_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
}
void launchLumpFactor(const DevSolver& d, double* hess, const int32_t* lumps, int nLumps, int maxRows, int maxWidth, int* errFlag, cudaStream_t s) {
  const size_t smem = (size_t)(maxWidth + kTrsmRows) * (maxWidth + 1) * sizeof(double);
  ensureSmem((const void*)lumpFactorKernel, smem);
  const int tiles = std::max(1, (maxRows + kTrsmRows - 1) / kTrsmRows);
  lumpFactorKernel<<<dim3(tiles, nLumps), 128, smem, s>>>(d, hess, lumps, errFlag);
  VB_CUDA(cudaGetLastError());
}
void launchLumpUpdate(const DevSolver& d, double* hess, const LevelTask* tasks, int nTasks, int maxWidth, int useAtomics, cudaStream_t s) {
  if (nTasks == 0) return;
  const size_t smem = (size_t)2 * kUpdTile * (maxWidth + 1) * sizeof(double);
  ensureSmem((const void*)lumpUpdateKernel, smem);
  lumpUpdateKernel<<<nTasks, 256, smem, s>>>(d, hess, tasks, useAtomics);
  VB_CUDA(cudaGetLastError());
}
void launchSolveLandmarksFwd(const DevSolver& d, const double* hess, double* vec, cudaStream_t s) {
  if (d.nP == 0) return;
  solveLmFwdKernel<<<(int)((d.nP * 32 + 255) / 256), 256, 0, s>>>(d, hess, vec);
  VB_CUDA(cudaGetLastError());
}
void launchSolveLandmarksBwd(const DevSolver& d, const double* hess, double* vec, cudaStream_t s) {
  if (d.nP == 0) return;
  solveLmBwdKernel<<<(int)((d.nP * 32 + 255) / 256), 256, 0, s>>>(d, hess, vec);
  VB_CUDA(cudaGetLastError());
}
void launchSolveLumpsFwd(const DevSolver& d, const double* hess, double* vec, const int32_t* lumps, int nLumps, int maxWidth, int useAtomics, cudaStream_t s) {
  solveLumpFwdKernel<<<nLumps, 128, maxWidth * sizeof(double), s>>>(d, hess, vec, lumps, useAtomics);
  VB_CUDA(cudaGetLastError());
}
void launchSolveLumpsBwd(const DevSolver& d, const double* hess, double* vec, const int32_t* lumps, int nLumps, int maxWidth, cudaStream_t s) {
  // dynamic smem = width + below rows of the widest lump of the problem (passed as maxWidth = w + r bound)
  const size_t smem = (size_t)maxWidth * sizeof(double);
  ensureSmem((const void*)solveLumpBwdKernel, smem);
  solveLumpBwdKernel<<<nLumps, 128, smem, s>>>(d, hess, vec, lumps);
  VB_CUDA(cudaGetLastError());
}

} // namespace vb
