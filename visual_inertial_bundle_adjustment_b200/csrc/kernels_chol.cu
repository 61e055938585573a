// Supernodal Cholesky of the reduced camera system and the supernode part of the
// solves (K11), level-scheduled over the elimination tree.  The dense contractions — panel trsm
// updates and the supernode outer product B B^T — run on the FP64 tensor pipe through
// mma.sync.aligned.m8n8k4.f64 (DMMA; tcgen05 has no f64 kind), operands staged in shared memory.
#include <map>
#include <mutex>
#include "vb_internal.hpp"
#include "dev_math.cuh"

namespace vb {

// shared-memory leading dimension: >= n rounded up to the 8-wide MMA tiles, == 4 mod 8 so that the
// 4-row x 4-column fragment loads of a half warp hit 16 distinct 8-byte banks
__host__ __device__ inline int ldFor(int n) { return ((n + 7) & ~7) + 4; }

// acc[mt][nt] (8x8 tiles) += A[rowBase + 8 mt .., :K] * B[colBase + 8 nt .., :K]^T, A/B row-major in smem
template <int MT, int NT>
VB_DEV void dmmaABt(const double* As, int lda, const double* Bs, int ldb, int K4, int rowBase, int colBase, double (&acc)[MT][NT][2]) {
  const int lane = threadIdx.x & 31;
  const double* ap = As + (rowBase + (lane >> 2)) * lda + (lane & 3);
  const double* bp = Bs + (colBase + (lane >> 2)) * ldb + (lane & 3);
  for (int k = 0; k < K4; k += 4) {
    double a[MT], b[NT];
#pragma unroll
    for (int m = 0; m < MT; m++) a[m] = ap[m * 8 * lda + k];
#pragma unroll
    for (int n = 0; n < NT; n++) b[n] = bp[n * 8 * ldb + k];
#pragma unroll
    for (int m = 0; m < MT; m++)
#pragma unroll
      for (int n = 0; n < NT; n++) dmma(acc[m][n][0], acc[m][n][1], a[m], b[n]);
  }
}

constexpr int kPanelRows = 64;  // panel rows per CTA in the factor kernel
constexpr int kNB = 8;          // column block of the blocked potrf / trsm

// ---- K10a: diagonal potrf (one CTA per supernode) and panel trsm (one CTA per 64 panel rows) -------
// Both are blocked by 8 columns: the bulk is [rows x 8] -= [rows x k] [8 x k]^T on DMMA, the 8x8
// triangles are solved per row in registers.  potrf publishes L into diagL (the panel's diagonal
// block in hess is left untouched); trsm reads it back into shared memory.
__global__ void __launch_bounds__(256) lumpPotrfKernel(DevSolver d, const double* hess, const int32_t* lumps, int* errFlag) {
  extern __shared__ double sm[];
  __shared__ int bad;
  __shared__ double dinv8[kNB];
  __shared__ double dinvAll[128];
  const int L = lumps[blockIdx.x];
  const int w = d.lumpWidth[L];
  const int ld = ldFor(w);
  const int wp = (w + 7) & ~7;          // rows padded to the 8-row MMA tiles
  double* A = sm;                        // wp x ld   (diagonal block -> L)
  const double* panel = hess + d.lumpOff[L];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) bad = 0;
  loadTileAsync(A, ld, wp, ld, panel, w, w, w); // the strict upper part is never read before it is overwritten
  cp_async_wait_all();
  __syncthreads();
  // ---- blocked left-looking Cholesky of A (lower), 8 columns at a time
  for (int j0 = 0; j0 < w; j0 += kNB) {
    const int nb = min(kNB, w - j0);
    // (1) A[j0:, j0:j0+8] -= A[j0:, :j0] A[j0:j0+8, :j0]^T   (DMMA, one 8-row tile per warp iteration)
    if (j0 > 0) {
      for (int rt = (j0 >> 3) + warp; rt * 8 < w; rt += 8) {
        double acc[1][1][2] = {{{0.0, 0.0}}};
        dmmaABt<1, 1>(A, ld, A, ld, j0, rt * 8, j0, acc);
        const int rr = rt * 8 + (lane >> 2), cc = j0 + 2 * (lane & 3);
        A[rr * ld + cc] -= acc[0][0][0];
        A[rr * ld + cc + 1] -= acc[0][0][1];
      }
    }
    __syncthreads();
    // (2) 8x8 diagonal block: Cholesky in registers by warp 0 (lane i < 8 owns row j0+i); one
    // rsqrt per pivot, columns exchanged with shuffles; reciprocals of the pivots go to dinv8
    if (warp == 0) {
      double a[kNB];
#pragma unroll
      for (int k = 0; k < kNB; k++) a[k] = lane < kNB ? A[(j0 + lane) * ld + j0 + k] : 0.0;
      bool ok = true;
#pragma unroll
      for (int j = 0; j < kNB; j++) {
        if (j < nb) {
          const double djj = __shfl_sync(0xffffffffu, a[j], j);
          if (!(djj > 0.0)) ok = false;
          const double rs = rsqrt(djj);
          const double lij = (lane == j) ? djj * rs : a[j] * rs;
          a[j] = lij;
          if (lane == j) { dinv8[j] = rs; dinvAll[j0 + j] = rs; }
#pragma unroll
          for (int k = j + 1; k < kNB; k++) {
            const double lkj = __shfl_sync(0xffffffffu, lij, k);
            if (lane >= k) a[k] = fma(-lij, lkj, a[k]);
          }
        }
      }
      if (!ok && lane == 0) bad = 1;
      if (lane < nb) {
#pragma unroll
        for (int k = 0; k < kNB; k++) A[(j0 + lane) * ld + j0 + k] = k <= lane ? a[k] : 0.0;
      }
    }
    __syncthreads();
    if (bad) { if (tid == 0) atomicExch(errFlag, 2); return; }
    // (3) rows below the diagonal block: A[i, j0:j0+8] <- A[i, j0:j0+8] L_jj^-T, one thread per row
    for (int i = j0 + nb + tid; i < w; i += 256) {
      double x[kNB];
#pragma unroll
      for (int c = 0; c < kNB; c++) x[c] = A[i * ld + j0 + c];
#pragma unroll
      for (int c = 0; c < kNB; c++) {
        if (c < nb) {
          double s = x[c];
#pragma unroll
          for (int k = 0; k < c; k++) s = fma(-x[k], A[(j0 + c) * ld + j0 + k], s);
          x[c] = s * dinv8[c];
        }
      }
#pragma unroll
      for (int c = 0; c < kNB; c++) if (c < nb) A[i * ld + j0 + c] = x[c];
    }
    __syncthreads();
  }
  double* out = d.diagL + d.diagLOff[L];
  for (int a = warp; a < w; a += 8)
    for (int b = lane; b < w; b += 32) out[a * w + b] = b <= a ? A[a * ld + b] : 0.0;
  // inverses of the 32x32 diagonal blocks of L (for the DMMA trsm): thread c owns column c
  double* inv = d.diagInv + (int64_t)d.camStart[d.lumpFirst[L]] * 32;
  if (tid < w) {
    // column-oriented forward substitution on e_c with the column in registers (compile-time indices
    // only: no local memory); the updates of one step are independent FMAs, L is read by broadcast
    const int c = tid, b0 = c & ~31, nbk = min(32, w - b0), cc = c - b0;
    const double* Lb = A + b0 * ld + b0;
    double z[32];
#pragma unroll
    for (int i = 0; i < 32; i++) z[i] = (i == cc) ? 1.0 : 0.0;
#pragma unroll
    for (int i = 0; i < 32; i++) {
      if (i < nbk) {
        const double zi = (i >= cc) ? z[i] * dinvAll[b0 + i] : 0.0;
        z[i] = zi;
#pragma unroll
        for (int k = i + 1; k < 32; k++) if (k < nbk) z[k] = fma(-zi, Lb[k * ld + i], z[k]);
      }
    }
#pragma unroll
    for (int i = 0; i < 32; i++) if (i < nbk) inv[(int64_t)(b0 + i) * 32 + cc] = z[i];
  }
}

// panel trsm X L^T = B by 32-column blocks: X_j <- (X_j - X_{<j} L_{j,<j}^T) Linv_jj^T, both products
// on DMMA.  Only the 32x32 diagonal blocks are inverted (by lumpPotrfKernel).  64 panel rows per CTA,
// one 8-row tile per warp (a warp only touches its own rows of X).
constexpr int kTB = 32;
// rhs != null: one more tile per lump carries the lump's segment of the right-hand side through the same
// substitution (x = b^T L^-T, i.e. the forward solve y_L = L^-1 b_L of Solver::solve falls out of the factorisation)
__global__ void __launch_bounds__(256) lumpTrsmKernel(DevSolver d, double* hess, const int32_t* lumps, double* rhs) {
  extern __shared__ double sm[];
  const int L = lumps[blockIdx.y];
  const int w = d.lumpWidth[L], r = d.lumpRows[L] - w;
  const bool rhsTile = rhs != nullptr && blockIdx.x == gridDim.x - 1;
  const int tile = blockIdx.x;
  if (!rhsTile && tile * kPanelRows >= r) return;
  const int nblk = (w + kTB - 1) / kTB;
  const int ld = ldFor(kTB * nblk);              // X (padded to whole 32-column blocks)
  const int ldl = ldFor(max(kTB * (nblk - 1), 8)); // L row block [32 x 32 jb]
  double* X = sm;                                 // 64 x ld
  double* Lr = X + (size_t)kPanelRows * ld;       // 32 x ldl
  double* Li = Lr + (size_t)kTB * ldl;            // 32 x 36
  double* panel = hess + d.lumpOff[L];
  const double* Ld = d.diagL + d.diagLOff[L];
  const double* inv = d.diagInv + (int64_t)d.camStart[d.lumpFirst[L]] * 32;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row0 = rhsTile ? 0 : tile * kPanelRows;
  const int nrows = rhsTile ? 1 : min(kPanelRows, r - row0);
  double* below = rhsTile ? rhs + d.camVec0 + d.camStart[d.lumpFirst[L]] : panel + (int64_t)(w + row0) * w;
  loadTileAsync(X, ld, kPanelRows, ld, below, w, nrows, w);
  for (int jb = 0; jb < nblk; jb++) {
    const int c0 = jb * kTB, nb = min(kTB, w - c0);
    __syncthreads();
    loadTileAsync(Lr, ldl, kTB, c0, Ld + (int64_t)c0 * w, w, nb, c0);
    loadTileAsync(Li, 36, kTB, 36, inv + (int64_t)c0 * 32, 32, nb, nb);
    cp_async_wait_all();
    __syncthreads();
    double acc[1][4][2];
#pragma unroll
    for (int n = 0; n < 4; n++) acc[0][n][0] = acc[0][n][1] = 0.0;
    if (jb > 0) dmmaABt<1, 4>(X, ld, Lr, ldl, c0, warp * 8, 0, acc);
    const int rr = warp * 8 + (lane >> 2);
#pragma unroll
    for (int n = 0; n < 4; n++) {
      const int cc = c0 + n * 8 + 2 * (lane & 3);
      X[rr * ld + cc] -= acc[0][n][0];
      X[rr * ld + cc + 1] -= acc[0][n][1];
      acc[0][n][0] = acc[0][n][1] = 0.0;
    }
    __syncwarp();
    dmmaABt<1, 4>(X + c0, ld, Li, 36, kTB, warp * 8, 0, acc);
    __syncwarp();
#pragma unroll
    for (int n = 0; n < 4; n++) {
      const int cc = c0 + n * 8 + 2 * (lane & 3);
      X[rr * ld + cc] = acc[0][n][0];
      X[rr * ld + cc + 1] = acc[0][n][1];
    }
  }
  __syncthreads();
  storeTile(below, w, X, ld, nrows, w);
}

// ---- K10b: supernode update  C = B B^T  (B = factored panel rows), scattered into the ancestors ----
// One CTA per 64x64 tile pair of the lump's below-diagonal rows; 8 warps, each a 16x32 sub-tile
// (2x4 DMMA tiles).  The scatter uses the per-lump relative-index tables built by the symbolic phase.
constexpr int kUpdTile = 64;
constexpr int kGemvSlice = 256;   // panel rows per fused forward-solve task (must match api.cu)
// Tasks with ti < 0 belong to the fused forward solve (rhs != null): slice tj of b_below -= P y_L, with y_L the
// lump's solved segment that lumpTrsmKernel left in rhs.
__global__ void __launch_bounds__(256) lumpUpdateKernel(DevSolver d, double* hess, const LevelTask* tasks, int useAtomics, double* rhs) {
  extern __shared__ double sm[];
  const LevelTask t = tasks[blockIdx.x];
  const int L = t.lump;
  const int w = d.lumpWidth[L], r = d.lumpRows[L] - w;
  if (t.ti < 0) {
    if (!rhs) return;
    double* x = sm;
    const double* vL = rhs + d.camVec0 + d.camStart[d.lumpFirst[L]];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < w; i += 256) x[i] = vL[i];
    __syncthreads();
    const double* below = hess + d.lumpOff[L] + (int64_t)w * w;
    const int64_t q0 = d.updPtr[L];
    const int rBeg = t.tj * kGemvSlice, rEnd = min(r, rBeg + kGemvSlice);
    for (int row = rBeg + warp * 2; row < rEnd; row += 16) {   // two rows per warp pass
      const double* P0 = below + (int64_t)row * w;
      const bool two = row + 1 < rEnd;
      const double* P1 = two ? P0 + w : P0;
      double s0 = 0, s1 = 0;
      for (int c = lane; c < w; c += 32) { const double xc = x[c]; s0 = fma(P0[c], xc, s0); s1 = fma(P1[c], xc, s1); }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) { s0 += __shfl_xor_sync(0xffffffffu, s0, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o); }
      if (lane == 0) {
        double* out = rhs + d.camVec0 + d.updGlobal[q0 + row];
        if (useAtomics) red_add(out, -s0); else *out -= s0;
      } else if (lane == 1 && two) {
        double* out = rhs + d.camVec0 + d.updGlobal[q0 + row + 1];
        if (useAtomics) red_add(out, -s1); else *out -= s1;
      }
    }
    return;
  }
  const int ld = ldFor(w);
  double* As = sm;
  double* Bs = (t.ti == t.tj) ? sm : sm + (size_t)kUpdTile * ld;
  const double* below = hess + d.lumpOff[L] + (int64_t)w * w;
  const int i0 = t.ti * kUpdTile, j0 = t.tj * kUpdTile;
  const int ni = min(kUpdTile, r - i0), nj = min(kUpdTile, r - j0);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  loadTileAsync(As, ld, kUpdTile, ld, below + (int64_t)i0 * w, w, ni, w);
  if (t.ti != t.tj) loadTileAsync(Bs, ld, kUpdTile, ld, below + (int64_t)j0 * w, w, nj, w);
  cp_async_wait_all();
  __syncthreads();
  const int rowBase = (warp >> 1) * 16, colBase = (warp & 1) * 32;
  double acc[2][4][2];
#pragma unroll
  for (int m = 0; m < 2; m++)
#pragma unroll
    for (int n = 0; n < 4; n++) acc[m][n][0] = acc[m][n][1] = 0.0;
  dmmaABt<2, 4>(As, ld, Bs, ld, (w + 3) & ~3, rowBase, colBase, acc);
  const int64_t q0 = d.updPtr[L];
  const int32_t* rel = d.rel + d.relPtr[L];
  // scatter in two phases so that the index loads of all 16 entries are in flight together (atomics
  // order memory operations, which would otherwise serialise every lookup behind the previous update)
  double* dst[4][2][2];
#pragma unroll
  for (int n = 0; n < 4; n++)
#pragma unroll
    for (int e = 0; e < 2; e++) {
      const int qj = j0 + colBase + n * 8 + 2 * (lane & 3) + e;
      const bool cv = qj < r;
      const int64_t colOff = cv ? d.updColOff[q0 + qj] : 0;
      const int stride = cv ? d.updStride[q0 + qj] : 0;
      const int32_t* relT = rel + (int64_t)(cv ? d.updTgt[q0 + qj] : 0) * r;
#pragma unroll
      for (int m = 0; m < 2; m++) {
        const int qi = i0 + rowBase + m * 8 + (lane >> 2);
        dst[n][e][m] = (cv && qi < r && qi >= qj) ? hess + colOff + (int64_t)relT[qi] * stride : nullptr;
      }
    }
#pragma unroll
  for (int n = 0; n < 4; n++)
#pragma unroll
    for (int e = 0; e < 2; e++)
#pragma unroll
      for (int m = 0; m < 2; m++) {
        double* p = dst[n][e][m];
        if (!p) continue;
        if (useAtomics) red_add(p, -acc[m][n][e]); else *p -= acc[m][n][e];
      }
}

// ---- K9b: Schur update of the camera system by landmark groups -------------------------------------
// A group is a few chunks of <= 32 landmarks whose row blocks have a small union (<= 1024 scalar
// rows).  Per chunk the dense operand U = [union rows] x [3 G] holds W_p L_p^-T of every member (zeros
// where a landmark does not see a block), gathered straight from the landmark panels; the update of
// the reduced system is -sum_chunks U U^T, accumulated tile by tile on DMMA (K loop over the chunks)
// and scattered ONCE with the same relative-index tables as the supernodes (source index nL + g):
// one reduction per target entry per group instead of one per landmark.
constexpr int kGrpMax = 32;
__global__ void __launch_bounds__(256) schurGroupUpdateKernel(DevSolver d, double* hess, const SchurTask* tasks) {
  extern __shared__ double sm[];
  const SchurTask t = tasks[blockIdx.x];
  const int g = t.group;
  const int r = d.grpRows[g];
  const int nch = d.grpChunks[g], nblk = d.grpBlks[g], nt = (r + kUpdTile - 1) / kUpdTile;
  constexpr int ld = 3 * kGrpMax + 4;       // 100
  double* As = sm;
  double* Bs = (t.ti == t.tj) ? sm : sm + (size_t)kUpdTile * ld;
  const int i0 = t.ti * kUpdTile, j0 = t.tj * kUpdTile;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t src = d.nL + g;
  const int64_t q0 = d.updPtr[src];
  const int64_t qg = q0 - d.grpRowBase;
  const uint32_t* tmask = d.grpTileMask + d.grpMaskPtr[g];
  const int rowBase = (warp >> 1) * 16, colBase = (warp & 1) * 32;
  double acc[2][4][2];
#pragma unroll
  for (int m = 0; m < 2; m++)
#pragma unroll
    for (int n = 0; n < 4; n++) acc[m][n][0] = acc[m][n][1] = 0.0;
  // this thread's gather row (4 threads per tile row) and its union block / sub-row, per side
  const int a = tid >> 2, kq = tid & 3;
  int rowBlk[2], rowSub3[2];
#pragma unroll
  for (int side = 0; side < 2; side++) {
    const int q = (side == 0 ? i0 : j0) + a;
    rowBlk[side] = q < r ? d.grpRowBlk[qg + q] : -1;
    rowSub3[side] = q < r ? 3 * d.grpRowSub[qg + q] : 0;
  }
  // K loop over the chunks of the group: gather [tile rows x 3 present members], accumulate on DMMA
  for (int c = 0; c < nch; c++) {
    const uint32_t mask = tmask[c * nt + t.ti] & tmask[c * nt + t.tj];
    if (!mask) continue;                     // uniform over the CTA
    const int cnt = __popc(mask);            // members of this chunk present in both row tiles
    const int K4 = (3 * cnt + 3) & ~3;
    const int32_t* blkOff = d.grpBlkOff + (d.grpBlkPtr[g] + (int64_t)c * nblk) * kGrpMax;
    __syncthreads();                         // the previous chunk's fragments have been read
    for (int side = 0; side < 2; side++) {
      if (side == 1 && t.ti == t.tj) break;
      double* o = (side == 0 ? As : Bs) + a * ld;
      if (kq == 0) for (int z = 3 * cnt; z < K4; z++) o[z] = 0.0;
      if (rowBlk[side] >= 0) {
        // thread kq of the row's quad owns member slots 8 kq .. 8 kq + 7: their 8 panel offsets come in
        // with two 16-byte loads (one round trip), then every present member's 3-double W row is copied
        // with cp.async into its compacted column (rank among the present members)
        const int4* bo = reinterpret_cast<const int4*>(blkOff + rowBlk[side] * kGrpMax + 8 * kq);
        const int4 o0 = bo[0], o1 = bo[1];
        const int offs[8] = {o0.x, o0.y, o0.z, o0.w, o1.x, o1.y, o1.z, o1.w};
        int k = __popc(mask & ((1u << (8 * kq)) - 1u));
#pragma unroll
        for (int z = 0; z < 8; z++) {
          if (!((mask >> (8 * kq + z)) & 1u)) continue;
          const int off = offs[z];
          const bool v = off >= 0;
          const double* w = v ? hess + off + rowSub3[side] : hess;
          cp_async8(o + 3 * k, w, v); cp_async8(o + 3 * k + 1, w + 1, v); cp_async8(o + 3 * k + 2, w + 2, v);
          k++;
        }
      } else {
        for (int k = kq; k < cnt; k += 4) { o[3 * k] = 0.0; o[3 * k + 1] = 0.0; o[3 * k + 2] = 0.0; }
      }
    }
    cp_async_wait_all();
    __syncthreads();
    dmmaABt<2, 4>(As, ld, Bs, ld, K4, rowBase, colBase, acc);
  }
  const int32_t* rel = d.rel + d.relPtr[src];
  // two-phase scatter: every index lookup first (independent loads in flight together), then the reductions.
  // The union rows are not in elimination order (most-shared rows first), so which of the two rows is the
  // column of the lower-triangular target is decided per entry from their reduced-system positions.
  double* dst[4][2][2];
  int64_t rowCol[2]; int rowStride[2], rowGlob[2]; const int32_t* rowRel[2];
#pragma unroll
  for (int m = 0; m < 2; m++) {
    const int qi = i0 + rowBase + m * 8 + (lane >> 2);
    const bool rv = qi < r;
    rowCol[m] = rv ? d.updColOff[q0 + qi] : 0;
    rowStride[m] = rv ? d.updStride[q0 + qi] : 0;
    rowGlob[m] = rv ? d.updGlobal[q0 + qi] : -1;
    rowRel[m] = rel + (int64_t)(rv ? d.updTgt[q0 + qi] : 0) * r;
  }
#pragma unroll
  for (int n = 0; n < 4; n++)
#pragma unroll
    for (int e = 0; e < 2; e++) {
      const int qj = j0 + colBase + n * 8 + 2 * (lane & 3) + e;
      const bool cv = qj < r;
      const int64_t colOff = cv ? d.updColOff[q0 + qj] : 0;
      const int stride = cv ? d.updStride[q0 + qj] : 0;
      const int gj = cv ? d.updGlobal[q0 + qj] : 0;
      const int32_t* relT = rel + (int64_t)(cv ? d.updTgt[q0 + qj] : 0) * r;
#pragma unroll
      for (int m = 0; m < 2; m++) {
        const int qi = i0 + rowBase + m * 8 + (lane >> 2);
        double* p = nullptr;
        if (cv && qi < r && qi >= qj) {
          if (rowGlob[m] >= gj) { const int rr = relT[qi]; if (rr >= 0) p = hess + colOff + (int64_t)rr * stride; }
          else { const int rr = rowRel[m][qj]; if (rr >= 0) p = hess + rowCol[m] + (int64_t)rr * rowStride[m]; }
        }
        dst[n][e][m] = p;
      }
    }
#pragma unroll
  for (int n = 0; n < 4; n++)
#pragma unroll
    for (int e = 0; e < 2; e++)
#pragma unroll
      for (int m = 0; m < 2; m++) {
        const double v = acc[m][n][e];
        if (dst[n][e][m] && v != 0.0) red_add(dst[n][e][m], -v);
      }
}

// ---- K11 (supernode part) ---------------------------------------------------------------------------
// Warp-level triangular solves: L (w x w, lower, row-major, odd leading dimension) and the
// reciprocals of its diagonal in shared memory, x in shared memory.  32 columns at a time live in
// registers (one per lane) and are eliminated with shuffles; earlier/later blocks enter as GEMVs.
VB_DEV void warpSolveL(const double* Ls, int ld, const double* dinv, double* x, int w) {  // L y = x
  const int lane = threadIdx.x & 31;
  for (int b0 = 0; b0 < w; b0 += 32) {
    const int k = b0 + lane;
    double xk = k < w ? x[k] : 0.0;
    if (k < w) for (int c = 0; c < b0; c++) xk = fma(-Ls[k * ld + c], x[c], xk);
    const int nb = min(32, w - b0);
    for (int c = 0; c < nb; c++) {
      const double xc = __shfl_sync(0xffffffffu, xk, c) * dinv[b0 + c];
      if (lane == c) xk = xc;
      else if (lane > c && k < w) xk = fma(-Ls[k * ld + b0 + c], xc, xk);
    }
    if (k < w) x[k] = xk;
    __syncwarp();
  }
}
VB_DEV void warpSolveLt(const double* Ls, int ld, const double* dinv, double* x, int w) { // L^T y = x
  const int lane = threadIdx.x & 31;
  for (int b0 = ((w - 1) / 32) * 32; b0 >= 0; b0 -= 32) {
    const int k = b0 + lane;
    double xk = k < w ? x[k] : 0.0;
    if (k < w) for (int c = b0 + 32; c < w; c++) xk = fma(-Ls[c * ld + k], x[c], xk);
    const int nb = min(32, w - b0);
    for (int c = nb - 1; c >= 0; c--) {
      const double xc = __shfl_sync(0xffffffffu, xk, c) * dinv[b0 + c];
      if (lane == c) xk = xc;
      else if (lane < c) xk = fma(-Ls[(b0 + c) * ld + k], xc, xk);
    }
    if (k < w) x[k] = xk;
    __syncwarp();
  }
}

constexpr int kSolveSlice = 128; // panel rows per CTA in the solves

// forward over the lumps of one level: x_L = L^-1 b_L ; b_below -= P x_L.
// grid = (row slices, lumps).  Every slice CTA solves the small triangular system redundantly and
// applies its 128 panel rows; the CTA that finishes last (per-lump counter) publishes x_L, which
// guarantees every other slice has already read b_L.
template <bool STAGE>
__global__ void __launch_bounds__(256) solveLumpFwdKernel(DevSolver d, const double* hess, double* vec, const int32_t* lumps, int useAtomics) {
  extern __shared__ double sm[];
  __shared__ int isLast;
  const int L = lumps[blockIdx.y];
  const int w = d.lumpWidth[L], r = d.lumpRows[L] - w;
  const int nSlices = max(1, (r + kSolveSlice - 1) / kSolveSlice);
  const int slice = blockIdx.x;
  if (slice >= nSlices) return;
  const int ld = w | 1;
  double* Ls = sm;
  double* x = sm + (size_t)w * ld;
  double* dinv = x + w;
  double* vL = vec + d.camVec0 + d.camStart[d.lumpFirst[L]];
  const double* Ld = d.diagL + d.diagLOff[L];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const double* below = hess + d.lumpOff[L] + (int64_t)w * w;
  const int rBeg = slice * kSolveSlice, rEnd = min(r, rBeg + kSolveSlice);
  double* Ps = dinv + w; // [kSolveSlice x w] staged panel rows (STAGE only)
  loadTileAsync(Ls, ld, w, w, Ld, w, w, w);
  if (STAGE && rEnd > rBeg) loadTileAsync(Ps, w, rEnd - rBeg, w, below + (int64_t)rBeg * w, w, rEnd - rBeg, w);
  for (int i = tid; i < w; i += 256) { x[i] = vL[i]; dinv[i] = 1.0 / Ld[i * w + i]; }
  cp_async_wait_all();
  __syncthreads();
  if (warp == 0) warpSolveL(Ls, ld, dinv, x, w);
  __syncthreads();
  const int64_t q0 = d.updPtr[L];
  for (int row = rBeg + warp * 2; row < rEnd; row += 16) { // two rows per warp pass
    const double* P0 = STAGE ? Ps + (row - rBeg) * w : below + (int64_t)row * w;
    const bool two = row + 1 < rEnd;
    const double* P1 = two ? P0 + w : P0;
    double s0 = 0, s1 = 0;
    for (int c = lane; c < w; c += 32) { const double xc = x[c]; s0 = fma(P0[c], xc, s0); s1 = fma(P1[c], xc, s1); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { s0 += __shfl_xor_sync(0xffffffffu, s0, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o); }
    if (lane == 0) {
      double* out = vec + d.camVec0 + d.updGlobal[q0 + row];
      if (useAtomics) red_add(out, -s0); else *out -= s0;
    } else if (lane == 1 && two) {
      double* out = vec + d.camVec0 + d.updGlobal[q0 + row + 1];
      if (useAtomics) red_add(out, -s1); else *out -= s1;
    }
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const int done = atomicAdd(d.solveCnt + L, 1);
    isLast = (done == nSlices - 1);
    if (isLast) d.solveCnt[L] = 0;
  }
  __syncthreads();
  if (isLast) for (int i = tid; i < w; i += 256) vL[i] = x[i];
}
// backward: x_L = L^-T (b_L - P^T x_below).  Slice CTAs accumulate their part of P^T x_below into a
// per-lump scratch with fp64 reductions; the last one to arrive finishes the triangular solve.
__global__ void __launch_bounds__(256) solveLumpBwdKernel(DevSolver d, const double* hess, double* vec, const int32_t* lumps) {
  extern __shared__ double sm[];
  __shared__ int isLast;
  const int L = lumps[blockIdx.y];
  const int w = d.lumpWidth[L], r = d.lumpRows[L] - w;
  const int nSlices = max(1, (r + kSolveSlice - 1) / kSolveSlice);
  const int slice = blockIdx.x;
  if (slice >= nSlices) return;
  const int ld = w | 1;
  double* x = sm;                    // w
  double* dinv = x + w;              // w
  double* part = dinv + w;           // 2 x w
  double* xb = part + 2 * w;         // kSolveSlice
  double* Ls = xb + kSolveSlice;     // max(w x ld, kSolveSlice x w): staged panel rows first, then L
  double* Ps = Ls;
  double* vL = vec + d.camVec0 + d.camStart[d.lumpFirst[L]];
  double* accL = d.solveAcc + d.camStart[d.lumpFirst[L]];
  const double* Ld = d.diagL + d.diagLOff[L];
  const double* below = hess + d.lumpOff[L] + (int64_t)w * w;
  const int64_t q0 = d.updPtr[L];
  const int tid = threadIdx.x;
  const int rBeg = slice * kSolveSlice, rEnd = min(r, rBeg + kSolveSlice);
  if (rEnd > rBeg) loadTileAsync(Ps, w, rEnd - rBeg, w, below + (int64_t)rBeg * w, w, rEnd - rBeg, w);
  for (int i = tid; i < rEnd - rBeg; i += 256) xb[i] = vec[d.camVec0 + d.updGlobal[q0 + rBeg + i]];
  cp_async_wait_all();
  __syncthreads();
  {
    const int half = tid >> 7, c = tid & 127;
    double s0 = 0, s1 = 0;
    if (c < w) {
      int row = half;
      for (; row + 2 < rEnd - rBeg; row += 4) { s0 = fma(Ps[row * w + c], xb[row], s0); s1 = fma(Ps[(row + 2) * w + c], xb[row + 2], s1); }
      for (; row < rEnd - rBeg; row += 2) s0 = fma(Ps[row * w + c], xb[row], s0);
      part[half * w + c] = s0 + s1;
    }
  }
  __syncthreads();
  if (nSlices > 1 && rEnd > rBeg) for (int c = tid; c < w; c += 256) red_add(accL + c, part[c] + part[w + c]);
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const int done = atomicAdd(d.solveCnt + L, 1);
    isLast = (done == nSlices - 1);
    if (isLast) d.solveCnt[L] = 0;
  }
  __syncthreads();
  if (!isLast) return;
  __threadfence();
  loadTileAsync(Ls, ld, w, w, Ld, w, w, w);
  for (int i = tid; i < w; i += 256) {
    double s;
    if (nSlices > 1) { s = __ldcg(accL + i); accL[i] = 0.0; } else s = part[i] + part[w + i];
    x[i] = vL[i] - s;
    dinv[i] = 1.0 / Ld[i * w + i];
  }
  cp_async_wait_all();
  __syncthreads();
  if (tid < 32) warpSolveLt(Ls, ld, dinv, x, w);
  __syncthreads();
  for (int i = tid; i < w; i += 256) vL[i] = x[i];
}

// ---- launchers ----------------------------------------------------------------------------------------
// The opt-in limit is a per-function, per-context attribute: raise it monotonically under a lock so
// that several host threads driving their own problems on one device never lower each other's limit.
static void ensureSmem(const void* fn, size_t bytes) {
  if (bytes <= 48 * 1024) return;
  static std::mutex mu;
  static std::map<std::pair<int, const void*>, size_t> cur;
  int dev = 0;
  VB_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  size_t& c = cur[{dev, fn}];
  if (bytes > c) {
    VB_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    c = bytes;
  }
}
void launchSchur(const DevSolver& d, double* hess, const SchurTask* tasks, int nTasks, int* errFlag, cudaStream_t s) {
  launchSchurPrep(d, hess, errFlag, s);
  launchSchurUpdate(d, hess, tasks, nTasks, s);
}
void launchSchurUpdate(const DevSolver& d, double* hess, const SchurTask* tasks, int nTasks, cudaStream_t s) {
  if (nTasks == 0) return;
  const size_t smem = (size_t)2 * kUpdTile * (3 * kGrpMax + 4) * sizeof(double);
  ensureSmem((const void*)schurGroupUpdateKernel, smem);
  schurGroupUpdateKernel<<<nTasks, 256, smem, s>>>(d, hess, tasks);
  VB_CUDA(cudaGetLastError());
}
void launchLumpFactor(const DevSolver& d, double* hess, const int32_t* lumps, int nLumps, int maxRows, int maxWidth, int* errFlag, cudaStream_t s, double* rhs) {
  const size_t smemP = (size_t)((maxWidth + 7) & ~7) * ldFor(maxWidth) * sizeof(double);
  ensureSmem((const void*)lumpPotrfKernel, smemP);
  lumpPotrfKernel<<<nLumps, 256, smemP, s>>>(d, hess, lumps, errFlag);
  VB_CUDA(cudaGetLastError());
  if (maxRows <= 0 && !rhs) return;
  const int nblk = (maxWidth + kTB - 1) / kTB;
  const size_t smemT = ((size_t)kPanelRows * ldFor(kTB * nblk) + (size_t)kTB * ldFor(std::max(kTB * (nblk - 1), 8)) + kTB * 36) * sizeof(double);
  ensureSmem((const void*)lumpTrsmKernel, smemT);
  lumpTrsmKernel<<<dim3((std::max(maxRows, 0) + kPanelRows - 1) / kPanelRows + (rhs ? 1 : 0), nLumps), 256, smemT, s>>>(d, hess, lumps, rhs);
  VB_CUDA(cudaGetLastError());
}
void launchLumpUpdate(const DevSolver& d, double* hess, const LevelTask* tasks, int nTasks, int maxWidth, int useAtomics, cudaStream_t s, double* rhs) {
  if (nTasks == 0) return;
  const size_t smem = (size_t)2 * kUpdTile * ldFor(maxWidth) * sizeof(double);
  ensureSmem((const void*)lumpUpdateKernel, smem);
  lumpUpdateKernel<<<nTasks, 256, smem, s>>>(d, hess, tasks, useAtomics, rhs);
  VB_CUDA(cudaGetLastError());
}
void launchSolveLumpsFwd(const DevSolver& d, const double* hess, double* vec, const int32_t* lumps, int nLumps, int maxWidth, int maxBelow, int useAtomics, cudaStream_t s) {
  const int slices = std::max(1, (maxBelow + kSolveSlice - 1) / kSolveSlice);
  const size_t base = ((size_t)maxWidth * (maxWidth | 1) + 2 * (size_t)maxWidth) * sizeof(double);
  const size_t staged = base + (size_t)kSolveSlice * maxWidth * sizeof(double);
  // narrow levels are latency bound: stage the panel slice with cp.async behind the triangular solve;
  // wide levels are throughput bound: keep the small footprint (2 CTAs per SM)
  if ((int64_t)slices * nLumps <= 2 * 148 && staged <= 200 * 1024) {
    ensureSmem((const void*)solveLumpFwdKernel<true>, staged);
    solveLumpFwdKernel<true><<<dim3(slices, nLumps), 256, staged, s>>>(d, hess, vec, lumps, useAtomics);
  } else {
    ensureSmem((const void*)solveLumpFwdKernel<false>, base);
    solveLumpFwdKernel<false><<<dim3(slices, nLumps), 256, base, s>>>(d, hess, vec, lumps, useAtomics);
  }
  VB_CUDA(cudaGetLastError());
}
void launchSolveLumpsBwd(const DevSolver& d, const double* hess, double* vec, const int32_t* lumps, int nLumps, int maxWidth, int maxBelow, cudaStream_t s) {
  const size_t tile = std::max((size_t)maxWidth * (maxWidth | 1), (size_t)kSolveSlice * maxWidth);
  const size_t smem = (tile + 4 * (size_t)maxWidth + kSolveSlice) * sizeof(double);
  ensureSmem((const void*)solveLumpBwdKernel, smem);
  const int slices = std::max(1, (maxBelow + kSolveSlice - 1) / kSolveSlice);
  solveLumpBwdKernel<<<dim3(slices, nLumps), 256, smem, s>>>(d, hess, vec, lumps);
  VB_CUDA(cudaGetLastError());
}

} // namespace vb
