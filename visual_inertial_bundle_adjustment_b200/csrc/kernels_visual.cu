// Visual-factor evaluation + assembly: FactorStore::computeGradHess for VisualFactor and
// RollingShutterVisualFactor (lib/small_thing/Factor.h:543-661,704-734; viba/problem/VisualFactor.cpp:40-82,131-210).
//
// Two kernels per store, each reading observation tables packed in its own processing order (DevStore::VisPack,
// coalesced loads, no index chains):
//   visualCameraKernel   rig-major: one CTA per chunk of <= 128 observations of one (rig pose, extrinsics,
//                        intrinsics).  Camera-side Jacobian rows, scaled by sqrt(rho'), are staged TRANSPOSED in
//                        shared memory (MT[column][row]: conflict-free 16-byte stores by the evaluating thread,
//                        conflict-free DMMA fragment loads) and contracted M^T M on the FP64 tensor pipe, the K
//                        range split over the four warps; one reduction per entry per chunk.
//                        Global shutter: J_pose = J_extr G with the chunk-constant G = [[R_cb, hat(t_cb) R_cb],
//                        [0, R_cb]] (the pose enters through T_cb T_bw), so only [J_extr | J_intr | r] (22 columns
//                        instead of 28) is staged and contracted, and the pose blocks follow in the epilogue as
//                        G^T S_EE G, S_EE G, S_CE G, G^T s_E.
//   visualLandmarkKernel landmark-major: thread = observation for the evaluation; the landmark x extrinsics /
//                        intrinsics blocks of a (landmark, calibration window) segment are then summed
//                        cooperatively: J rows transposed through a per-warp shared-memory tile, lane = output
//                        entry, one pass over the 32 observations, one reduction per entry per segment.
#include "vb_internal.hpp"
#include "dev_factors_ext.cuh"

namespace vb {

constexpr int kVisChunk2 = 128;

template <int N>
__device__ void blockReduceAddV(double (&v)[N], double* out) {
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

// ---- camera side -------------------------------------------------------------------------------------------
template <bool RS>
__global__ void __launch_bounds__(128, RS ? 2 : 4) visualCameraKernel(DevStore st, DevVars V, double* grad, double* hess, int updateCache,
                                                                        int dontRetry, double* costAcc) {
  constexpr int NV = RS ? 5 : 4, NP = NV * (NV + 1) / 2;
  constexpr int CI = RS ? 17 : 15;
  constexpr int NS = RS ? (12 + CI + 3 + 1) : (6 + CI + 1);  // staged columns: RS [T6 E6 C17 V3 r] = 33, GS [E6 C15 r] = 22
  constexpr int NT = (NS + 7) / 8;                            // 8-row tiles of the staged columns: 5 / 3
  constexpr int NQ = NT * (NT + 1) / 2;                       // lower tile pairs: 15 / 6
  constexpr int LDT = 2 * kVisChunk2 + 4;                     // 260 == 4 (mod 8)
  constexpr int NC = RS ? 32 : 27;                            // logical camera columns [T E C (V)]; the residual is index NC
  constexpr int LS = NS + 1, LL = NC + 2;                     // leading dimensions of S and L
  extern __shared__ double sm[];
  double* MT = sm;                                            // [NT*8][LDT]
  __shared__ double camv[40];                                 // pose(7) extr(7) intr(17) vel(3)
  __shared__ double Gm[36];
  const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int s0 = st.chunkStart[c], n = st.chunkStart[c + 1] - s0;
  const int4 h0 = st.pk.chunkHdr[2 * c], h1 = st.pk.chunkHdr[2 * c + 1];  // {pose, extr, intr, vel} data offsets, {flags, tdim, f0, -}
  {
    if (tid < 7) camv[tid] = V.data[h0.x + tid];
    else if (tid < 14) camv[tid] = V.data[h0.y + tid - 7];
    else if (tid < 31) camv[tid] = V.data[h0.z + tid - 14];
    else if (RS && tid < 34) camv[tid] = V.data[h0.w + tid - 31];
  }
  const uint32_t camFlags = (uint32_t)h1.x;
  const int camDim = h1.y, f0 = h1.z;
  __syncthreads();
  if (!RS && tid < 36) {  // G = [[R_cb, hat(t_cb) R_cb], [0, R_cb]]
    double R[9];
    q2R(camv + 7, R);
    const double* t = camv + 11;
    const int i = tid / 6, j = tid % 6;
    double g = 0.0;
    if (i < 3 && j < 3) g = R[i * 3 + j];
    else if (i >= 3 && j >= 3) g = R[(i - 3) * 3 + (j - 3)];
    else if (i < 3 && j >= 3) {  // (hat(t) R)[i][j-3]
      const int jj = j - 3;
      const double h[3][3] = {{0, -t[2], t[1]}, {t[2], 0, -t[0]}, {-t[1], t[0], 0}};
      g = h[i][0] * R[jj] + h[i][1] * R[3 + jj] + h[i][2] * R[6 + jj];
    }
    Gm[tid] = g;
  }
  double acc[1] = {0.0};
  {
    double2* col = reinterpret_cast<double2*>(MT + 2 * tid);  // column j of this thread's two rows: col[j * LDT / 2]
    bool live = false;
    if (tid < n) {
      const int f = st.order[s0 + tid];
      if (!(dontRetry && st.cache[f] < 0.0)) {
        double r[2], JX[6], JT[12], JE[12], JC[2 * CI], JV[6];
        const double* X = V.data + st.pk.cLm[s0 + tid];
        const double* cst = st.pk.cConst + (int64_t)(s0 + tid) * st.cpf;
        double cl[8];
#pragma unroll
        for (int j = 0; j < (RS ? 8 : 6); j++) cl[j] = cst[j];
        int status;
        if (RS) status = rs_visual_factor<true>(X, camv, camv + 7, camv + 14, camFlags, camv + 31, cl, st.aux, r, JX, JT, JE, JC, JV);
        else status = visual_factor<true>(X, camv, camv + 7, camv + 14, cl, r, JX, JT, JE, JC) ? 0 : 1;
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
          constexpr int H = LDT / 2;
          if (RS) {
#pragma unroll
            for (int j = 0; j < 6; j++) { col[j * H] = make_double2(sw * JT[j], sw * JT[6 + j]); col[(6 + j) * H] = make_double2(sw * JE[j], sw * JE[6 + j]); }
#pragma unroll
            for (int j = 0; j < CI; j++) col[(12 + j) * H] = make_double2(sw * JC[j], sw * JC[CI + j]);
#pragma unroll
            for (int j = 0; j < 3; j++) col[(12 + CI + j) * H] = make_double2(sw * JV[j], sw * JV[3 + j]);
          } else {
#pragma unroll
            for (int j = 0; j < 6; j++) col[j * H] = make_double2(sw * JE[j], sw * JE[6 + j]);
#pragma unroll
            for (int j = 0; j < CI; j++) col[(6 + j) * H] = make_double2(sw * JC[j], sw * JC[CI + j]);
          }
          col[(NS - 1) * H] = make_double2(sw * r[0], sw * r[1]);
        }
      }
    }
    constexpr int H = LDT / 2;
    if (!live) {
#pragma unroll
      for (int j = 0; j < NS; j++) col[j * H] = make_double2(0.0, 0.0);
    }
#pragma unroll
    for (int j = NS; j < NT * 8; j++) col[j * H] = make_double2(0.0, 0.0);
  }
  __syncthreads();
  // M^T M on DMMA: warp w contracts rows [64 w, 64 w + 64) for every lower tile pair
  double C[NQ][2];
#pragma unroll
  for (int q = 0; q < NQ; q++) C[q][0] = C[q][1] = 0.0;
  if (64 * warp < 2 * n) {
    const double* fp = MT + (lane >> 2) * LDT + 64 * warp + (lane & 3);
#pragma unroll 4
    for (int k = 0; k < 64; k += 4) {
      double fr[NT];
#pragma unroll
      for (int t = 0; t < NT; t++) fr[t] = fp[t * 8 * LDT + k];
      int q = 0;
#pragma unroll
      for (int ti = 0; ti < NT; ti++)
#pragma unroll
        for (int tj = 0; tj <= ti; tj++) { dmma(C[q][0], C[q][1], fr[ti], fr[tj]); q++; }
    }
  }
  __syncthreads();  // every warp is done reading MT: reuse it for the partial tiles
  double* part = sm;                       // [4][NQ][64]
  double* S = sm + 4 * NQ * 64;            // [NS][LS]  symmetric sum over the chunk
  double* L = S + NS * LS;                 // [NC+1][LL] logical [T E C (V) r] matrix (lower part used)
  {
    double* pw = part + warp * NQ * 64 + (lane >> 2) * 8 + 2 * (lane & 3);
#pragma unroll
    for (int q = 0; q < NQ; q++) { pw[q * 64] = C[q][0]; pw[q * 64 + 1] = C[q][1]; }
  }
  __syncthreads();
  for (int e = tid; e < NS * NS; e += 128) {
    const int i = e / NS, j = e % NS;
    if (j > i) continue;
    const int ti = i >> 3, tj = j >> 3, q = ti * (ti + 1) / 2 + tj;
    const double* p = part + q * 64 + (i & 7) * 8 + (j & 7);
    const double v = p[0] + p[NQ * 64] + p[2 * NQ * 64] + p[3 * NQ * 64];
    S[i * LS + j] = v; S[j * LS + i] = v;
  }
  __syncthreads();
  if (RS) {
    for (int e = tid; e < NS * NS; e += 128) { const int i = e / NS, j = e % NS; L[i * LL + j] = S[i * LS + j]; }
  } else {
    // S: E 0-5, C 6-20, r 21.  L: T 0-5, E 6-11, C 12-26, r 27
    // P = S_EE G (6x6), Q = S_CE G (15x6), sE G (1x6): rows 0..21 of S (E, C, r) times G
    for (int e = tid; e < 22 * 6; e += 128) {
      const int i = e / 6, j = e % 6;
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < 6; k++) s = fma(S[i * LS + k], Gm[k * 6 + j], s);
      L[(6 + i) * LL + j] = s;  // rows E (6-11), C (12-26), r (27); columns T
    }
    for (int e = tid; e < 22 * 22; e += 128) {  // the E / C / r part carries over
      const int i = e / 22, j = e % 22;
      L[(6 + i) * LL + 6 + j] = S[i * LS + j];
    }
    __syncthreads();
    if (tid < 36) {  // T x T = G^T P
      const int i = tid / 6, j = tid % 6;
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < 6; k++) s = fma(Gm[k * 6 + i], L[(6 + k) * LL + j], s);
      L[i * LL + j] = s;
    }
  }
  __syncthreads();
  // scatter: entry (i, j <= i) of the camera-side Hessian, and the gradient from row NC
  {
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
    const int64_t* gO = st.gOff + (int64_t)f0 * NV;
    const int64_t* hO = st.hOff + (int64_t)f0 * NP;
    const int32_t* hS = st.hStride + (int64_t)f0 * NP;
    for (int i = tid; i < NC; i += 128) {
      int bi, ai;
      locate(i, bi, ai);
      const int64_t gi = gO[1 + bi];
      const double v = L[NC * LL + i];
      if (gi >= 0 && ai >= 0 && v != 0.0) red_add(grad + gi + ai, v);
    }
    if (hess) {
      for (int e = tid; e < NC * NC; e += 128) {
        const int i = e / NC, j = e % NC;
        if (j > i) continue;
        const double v = L[i * LL + j];
        if (v == 0.0) continue;
        int bi, ai, bj, aj;
        locate(i, bi, ai);
        locate(j, bj, aj);
        if (ai < 0 || aj < 0 || gO[1 + bi] < 0) continue;
        const int pi = (bi + 1) * (bi + 2) / 2 + (bj + 1);
        const int64_t ho = hO[pi];
        if (ho < 0) continue;
        const int sw_ = hS[pi];
        const int stride = sw_ & 0x3fffffff;
        const bool tr = (sw_ >> 30) & 1;
        red_add(hess + ho + (tr ? (int64_t)aj * stride + ai : (int64_t)ai * stride + aj), v);
      }
    }
  }
  blockReduceAddV<1>(acc, costAcc);
}

// ---- landmark side -------------------------------------------------------------------------------------------
template <int N>
VB_DEV void segReduceV(double (&v)[N], int lane, int segEnd) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
    for (int i = 0; i < N; i++) {
      const double t = __shfl_down_sync(0xffffffffu, v[i], d);
      if (lane + d <= segEnd) v[i] += t;
    }
  }
}
VB_DEV int segEndOfV(uint32_t heads, int lane) {
  const uint32_t higher = lane == 31 ? 0u : (heads & ~((2u << lane) - 1u));
  return higher ? __ffs(higher) - 2 : 31;
}

// Observations in landmark-major order (sorted by landmark, extrinsics, intrinsics).  lOut per observation:
// {g_p offset, H_pp, W_pose, W_extr, W_intr, W_vel} as 32-bit offsets into grad / hess (0xffffffff = absent).
template <bool RS>
__global__ void __launch_bounds__(128, RS ? 3 : 4) visualLandmarkKernel(DevStore st, DevVars V, double* grad, double* hess, int dontRetry) {
  constexpr int CI = RS ? 17 : 15;
  constexpr int NE = 6 + CI;            // rows of [W_extr ; W_intr]: 21 / 23
  constexpr int NCOL = 2 * NE + 6;      // per-warp tile: JE, JC of residual row 0, of row 1, then w J_X (2 x 3)
  constexpr int LD = 33;
  constexpr uint32_t kNone = 0xffffffffu;
  extern __shared__ double sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* T = sm + warp * (NCOL * LD);
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const bool inRange = i < st.nActive;
  const int64_t ii = inRange ? i : st.nActive - 1;
  const int32_t* lv = st.pk.lVar + ii * 6;            // {X, pose, extr, intr, vel, camFlags}
  const int vX = inRange ? lv[0] : -1 - lane, vP = lv[1], vE = inRange ? lv[2] : -1, vC = inRange ? lv[3] : -1;
  const uint32_t* lo = st.pk.lOut + ii * 6;
  const uint32_t oG = lo[0], oPP = lo[1], oPose = lo[2], oExtr = lo[3], oIntr = lo[4], oVel = lo[5];
  // segment heads: landmark segments, and (landmark, extrinsics, intrinsics) segments inside them
  const int pL = __shfl_up_sync(0xffffffffu, vX, 1), pE = __shfl_up_sync(0xffffffffu, vE, 1), pC = __shfl_up_sync(0xffffffffu, vC, 1);
  const bool headL = lane == 0 || pL != vX;
  const bool headC = headL || pE != vE || pC != vC;
  const uint32_t headsC = __ballot_sync(0xffffffffu, headC);
  const int endL = segEndOfV(__ballot_sync(0xffffffffu, headL), lane);
  const uint32_t endsC = (headsC >> 1) | 0x80000000u;   // lane o ends a calibration segment
  double r[2] = {0, 0}, JX[6], JT[12], JE[12], JC[2 * CI], JV[6], X0[3] = {0, 0, 0}, X1[3] = {0, 0, 0};
  const uint32_t camFlags = RS ? (uint32_t)lv[5] : 0u;
  bool live = false;
  if (inRange && !(dontRetry && st.cache[st.orderLm[i]] < 0.0)) {
    const double* cst = st.pk.lConst + i * st.cpf;
    double cl[8];
#pragma unroll
    for (int j = 0; j < (RS ? 8 : 6); j++) cl[j] = cst[j];
    const double* X = V.data + vX;
    const double* Tbw = V.data + vP;
    const double* Tcb = V.data + vE;
    const double* cam = V.data + vC;
    int status;
    if (RS) status = rs_visual_factor<true>(X, Tbw, Tcb, cam, camFlags, V.data + lv[4], cl, st.aux, r, JX, JT, JE, JC, JV);
    else status = visual_factor<true>(X, Tbw, Tcb, cam, cl, r, JX, JT, JE, JC) ? 0 : 1;
    live = status == 0;
  }
  if (live) {
    double soft, w;
    loss_jet2(st.loss_a, st.loss_k, r[0] * r[0] + r[1] * r[1], soft, w);
#pragma unroll
    for (int a = 0; a < 3; a++) { X0[a] = w * JX[a]; X1[a] = w * JX[3 + a]; }
#pragma unroll
    for (int s = 0; s < 6; s++) { T[s * LD + lane] = JE[s]; T[(NE + s) * LD + lane] = JE[6 + s]; }
#pragma unroll
    for (int s = 0; s < CI; s++) { T[(6 + s) * LD + lane] = JC[s]; T[(NE + 6 + s) * LD + lane] = JC[CI + s]; }
  } else {
#pragma unroll
    for (int s = 0; s < 2 * NE; s++) T[s * LD + lane] = 0.0;
#pragma unroll
    for (int j = 0; j < 6; j++) JX[j] = 0.0;
  }
#pragma unroll
  for (int a = 0; a < 3; a++) { T[(2 * NE + a) * LD + lane] = X0[a]; T[(2 * NE + 3 + a) * LD + lane] = X1[a]; }
  const bool wr = oG != kNone;
  // gradient + H_pp: one write per landmark segment
  {
    double g[3];
#pragma unroll
    for (int a = 0; a < 3; a++) g[a] = r[0] * X0[a] + r[1] * X1[a];
    segReduceV<3>(g, lane, endL);
    if (headL && wr) {
#pragma unroll
      for (int a = 0; a < 3; a++) if (g[a] != 0.0) red_add(grad + oG + a, g[a]);
    }
    if (hess) {
      double h[6];
      int q = 0;
#pragma unroll
      for (int a = 0; a < 3; a++)
#pragma unroll
        for (int b = 0; b <= a; b++) h[q++] = X0[a] * JX[b] + X1[a] * JX[3 + b];
      segReduceV<6>(h, lane, endL);
      if (headL && wr && oPP != kNone) {
        double* Hpp = hess + oPP;
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
    if (oPose != kNone) {
      double* W = hess + oPose;
#pragma unroll
      for (int s_ = 0; s_ < 6; s_++)
#pragma unroll
        for (int a = 0; a < 3; a++) red_add(W + s_ * 3 + a, JT[s_] * X0[a] + JT[6 + s_] * X1[a]);
    }
    if (RS && oVel != kNone) {
      double* W = hess + oVel;
#pragma unroll
      for (int s_ = 0; s_ < 3; s_++)
#pragma unroll
        for (int a = 0; a < 3; a++) red_add(W + s_ * 3 + a, JV[s_] * X0[a] + JV[3 + s_] * X1[a]);
    }
  }
  __syncwarp();
  // landmark x extrinsics / intrinsics: lane = entry (row s of [W_extr ; W_intr], landmark coordinate a); one pass over
  // the 32 observations of the tile, flushed at every segment end
  constexpr int NENT = 3 * NE;
#pragma unroll
  for (int round = 0; round < (NENT + 31) / 32; round++) {
    const int entry = lane + 32 * round;
    const bool ev = entry < NENT;
    const int s = ev ? entry / 3 : 0, a = ev ? entry % 3 : 0;
    const double* j0 = T + s * LD;
    const double* j1 = T + (NE + s) * LD;
    const double* x0 = T + (2 * NE + a) * LD;
    const double* x1 = T + (2 * NE + 3 + a) * LD;
    // target row inside the block (RS: fixed columns 15/16 map to the tangent layout [15, readout?, time offset?])
    int row = s < 6 ? s : s - 6;
    bool on = ev;
    if (RS && s - 6 == 15) on = on && (camFlags & VB_CAM_EST_READOUT);
    if (RS && s - 6 == 16) { on = on && (camFlags & VB_CAM_EST_TIMEOFFSET); row = (camFlags & VB_CAM_EST_READOUT) ? 16 : 15; }
    double accv = 0.0;
#pragma unroll 8
    for (int o = 0; o < 32; o++) {
      accv = fma(j0[o], x0[o], accv);
      accv = fma(j1[o], x1[o], accv);
      if ((endsC >> o) & 1u) {   // warp-uniform
        const uint32_t tE = __shfl_sync(0xffffffffu, oExtr, o), tC = __shfl_sync(0xffffffffu, oIntr, o);
        const uint32_t tgt = s < 6 ? tE : tC;
        const uint32_t tg = __shfl_sync(0xffffffffu, oG, o);
        if (on && tgt != kNone && tg != kNone && accv != 0.0) red_add(hess + tgt + row * 3 + a, accv);
        accv = 0.0;
      }
    }
  }
}

// ---- refinePoints -------------------------------------------------------------------------------
// viba/problem/PointRefinement.cpp:96-196: a 3x3 Levenberg-Marquardt per landmark (<= 5 iterations) over its visual
// factors only — pointGradHess / pointCost (:49-86) through FactorStore::varGradHess (Factor.h:419-468) and singleCost
// (:230-232).  One warp per landmark: lanes take the landmark's observations (landmark-major tables of VisPack, store
// by store), cost / gradient / 3x3 Hessian are summed with shuffles, and every lane takes the same decisions.
// backups = the per-observation costBackups of the reference (a per-store scratch array).
VB_DEV bool solve3v(const double* H, const double* g, double* x) {   // H x = g, H SPD (lower entries used)
  if (!(H[0] > 0)) return false;
  const double l00 = sqrt(H[0]);
  const double l10 = H[3] / l00, l20 = H[6] / l00;
  const double d1 = H[4] - l10 * l10;
  if (!(d1 > 0)) return false;
  const double l11 = sqrt(d1), l21 = (H[7] - l20 * l10) / l11;
  const double d2 = H[8] - l20 * l20 - l21 * l21;
  if (!(d2 > 0)) return false;
  const double l22 = sqrt(d2);
  const double y0 = g[0] / l00, y1 = (g[1] - l10 * y0) / l11, y2 = (g[2] - l20 * y0 - l21 * y1) / l22;
  x[2] = y2 / l22;
  x[1] = (y1 - l21 * x[2]) / l11;
  x[0] = (y0 - l10 * x[1] - l20 * x[2]) / l00;
  return true;
}
VB_DEV double warpSum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// residual (+ J_X) of observation i of store S at landmark position X
template <bool WANT_J>
VB_DEV int refineEval(const RefineStore& S, const DevVars& V, int64_t i, const double* X, double* r, double* JX) {
  const int32_t* lv = S.lVar + i * 6;
  const double* cst = S.lConst + i * S.cpf;
  double JT[12], JE[12], JC[34], JV[6];
  if (S.rs) {
    double cl[8];
#pragma unroll
    for (int j = 0; j < 8; j++) cl[j] = cst[j];
    return rs_visual_factor<WANT_J>(X, V.data + lv[1], V.data + lv[2], V.data + lv[3], (uint32_t)lv[5], V.data + lv[4], cl, S.aux, r, JX, JT, JE, JC, JV);
  }
  double cl[6];
#pragma unroll
  for (int j = 0; j < 6; j++) cl[j] = cst[j];
  return visual_factor<WANT_J>(X, V.data + lv[1], V.data + lv[2], V.data + lv[3], cl, r, JX, JT, JE, JC) ? 0 : 1;
}

__global__ void __launch_bounds__(128) refinePointsKernel(RefineArgs A, DevVars V, double* perPoint, double* statsAcc) {
  constexpr double kLambda = 1e-5, kCostTolerance = 1e-8, kStepTolerance = 1e-6, kMinImprovementRatio = 0.2, kStepReduction = 0.3;
  constexpr int kMaxNumIterations = 5;
  const int lane = threadIdx.x & 31;
  const int64_t j = blockIdx.x * 4ll + (threadIdx.x >> 5);
  if (j >= A.nLm) return;
  double* Xg = V.data + A.lmOff[j];
  double X[3] = {Xg[0], Xg[1], Xg[2]};
  // (cost, grad, hess) at X; update: write the backups
  auto gradHess = [&](bool update, double* grad, double* hess) {
    double acc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};   // cost, g(3), lower H (6)
    for (int s = 0; s < A.nStores; s++) {
      const RefineStore& S = A.st[s];
      for (int64_t i = S.lmPtr[j] + lane; i < S.lmPtr[j + 1]; i += 32) {
        double r[2], JX[6];
        const int status = refineEval<true>(S, V, i, X, r, JX);
        if (status == 2) V.errAcc[0] = 1.0;
        if (status != 0) { if (update) S.backup[i] = -1.0; continue; }
        double soft, d;
        loss_jet2(S.loss_a, S.loss_k, r[0] * r[0] + r[1] * r[1], soft, d);
        acc[0] += 0.5 * soft;
        if (update) S.backup[i] = 0.5 * soft;
        int q = 4;
#pragma unroll
        for (int a = 0; a < 3; a++) {
          acc[1 + a] += d * (r[0] * JX[a] + r[1] * JX[3 + a]);
#pragma unroll
          for (int b = 0; b <= a; b++) acc[q++] += d * (JX[a] * JX[b] + JX[3 + a] * JX[3 + b]);
        }
      }
    }
#pragma unroll
    for (int q = 0; q < 10; q++) acc[q] = warpSum(acc[q]);
    grad[0] = acc[1]; grad[1] = acc[2]; grad[2] = acc[3];
    hess[0] = acc[4]; hess[3] = hess[1] = acc[5]; hess[4] = acc[6]; hess[6] = hess[2] = acc[7]; hess[7] = hess[5] = acc[8]; hess[8] = acc[9];
    return acc[0];
  };
  auto pointCost = [&]() {
    double c = 0.0;
    for (int s = 0; s < A.nStores; s++) {
      const RefineStore& S = A.st[s];
      for (int64_t i = S.lmPtr[j] + lane; i < S.lmPtr[j + 1]; i += 32) {
        const double bk = S.backup[i];
        if (bk < 0) continue;
        double r[2];
        const int status = refineEval<false>(S, V, i, X, r, nullptr);
        if (status != 0) { c += bk; continue; }
        double soft, d;
        loss_jet2(S.loss_a, S.loss_k, r[0] * r[0] + r[1] * r[1], soft, d);
        c += 0.5 * soft;
      }
    }
    return warpSum(c);
  };
  double startCost = 0, endCost = 0;
  int nIts = 0;
  for (int it = 0; it < kMaxNumIterations; it++) {
    double grad[3], hess[9];
    const double cost = gradHess(true, grad, hess);
    __syncwarp();   // the backups written above are read by other lanes' loops only at the same indices, but keep the order explicit
    if (it == 0) startCost = endCost = cost;
#pragma unroll
    for (int a = 0; a < 3; a++) hess[a * 4] = hess[a * 4] * (1.0 + kLambda) + kLambda;
    double step[3];
    if (!solve3v(hess, grad, step)) { nIts = -1; break; }
#pragma unroll
    for (int a = 0; a < 3; a++) step[a] = -step[a];
    const double bak[3] = {X[0], X[1], X[2]};
    bool success = false;
#pragma unroll
    for (int a = 0; a < 3; a++) X[a] += step[a];
    double newCost = pointCost();
    const double modelCostDelta = step[0] * grad[0] + step[1] * grad[1] + step[2] * grad[2];
    if (-modelCostDelta < kCostTolerance) break;   // the step stays applied, as in the reference
    if (newCost < cost + modelCostDelta * kMinImprovementRatio) {
      success = true;
    } else {
      double nGrad[3], nHess[9];
      gradHess(false, nGrad, nHess);
#pragma unroll
      for (int a = 0; a < 3; a++) X[a] = bak[a];
      const double newDelta = step[0] * nGrad[0] + step[1] * nGrad[1] + step[2] * nGrad[2];
      const double f = newDelta > 0 ? -modelCostDelta / (newDelta - modelCostDelta) : kStepReduction;
#pragma unroll
      for (int a = 0; a < 3; a++) { step[a] *= f; X[a] += step[a]; }
      newCost = pointCost();
      if (newCost < cost + (step[0] * grad[0] + step[1] * grad[1] + step[2] * grad[2]) * kMinImprovementRatio) success = true;
      else {
#pragma unroll
        for (int a = 0; a < 3; a++) X[a] = bak[a];
      }
    }
    if (success) { nIts++; endCost = newCost; }
    else { nIts = -1; break; }
    if (step[0] * step[0] + step[1] * step[1] + step[2] * step[2] < kStepTolerance * kStepTolerance) break;
  }
  if (lane == 0) {
    Xg[0] = X[0]; Xg[1] = X[1]; Xg[2] = X[2];
    if (perPoint) { perPoint[3 * j] = startCost; perPoint[3 * j + 1] = endCost; perPoint[3 * j + 2] = (double)nIts; }
    red_add(statsAcc + 0, startCost); red_add(statsAcc + 1, endCost); red_add(statsAcc + 2, 1.0);
    if (nIts < 0) red_add(statsAcc + 3, 1.0);
    else { if (nIts > 0) { red_add(statsAcc + 4, (double)nIts); red_add(statsAcc + 5, 1.0); } }
  }
}

void launchRefinePoints(const RefineArgs& A, const DevVars& v, double* perPoint, double* statsAcc, cudaStream_t s) {
  if (A.nLm <= 0) return;
  refinePointsKernel<<<(unsigned)((A.nLm + 3) / 4), 128, 0, s>>>(A, v, perPoint, statsAcc);
  VB_CUDA(cudaGetLastError());
}

// ---- refineTriangulationResult (second half) -------------------------------------------------------
// viba/single_session/Triangulation.cpp:99-161, the refinement passes of triangulatePoint (:219-238), for every landmark at
// once: maxIts undamped Gauss-Newton steps on the point with HuberLoss(radius) (SoftLoss.h:64-113) on the weighted
// reprojection error; each iteration re-classifies the observations (inlier = squared unweighted image error < thr2,
// :139-144; the unweighted error is sqrtH^-1 times the factor's residual) and with skipOutliers the outliers stay out of
// the normal equations.  Same warp-per-landmark layout and the same refineEval as refinePointsKernel above.  The inlier
// flags of the last iteration are left in the stores' per-observation scratch (RefineStore::backup: 1 inlier, 0 not);
// perPoint = {numInliers, cost of the last iteration, status (0, or -1: 3x3 system not positive definite, point kept)}.
__global__ void __launch_bounds__(128) refineTriangulationKernel(RefineArgs A, DevVars V, double thr2, int skipOutliers, int maxIts, double radius,
                                                                 double* perPoint, double* statsAcc) {
  const int lane = threadIdx.x & 31;
  const int64_t j = blockIdx.x * 4ll + (threadIdx.x >> 5);
  if (j >= A.nLm) return;
  double* Xg = V.data + A.lmOff[j];
  double X[3] = {Xg[0], Xg[1], Xg[2]};
  const double b = radius * radius;
  double nIn = 0.0, cost = 0.0;
  int status = 0;
  for (int it = 0; it < maxIts; it++) {
    double acc[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};   // cost, g(3), lower H (6), inliers
    for (int s = 0; s < A.nStores; s++) {
      const RefineStore& S = A.st[s];
      for (int64_t i = S.lmPtr[j] + lane; i < S.lmPtr[j + 1]; i += 32) {
        double r[2], JX[6];
        const int st = refineEval<true>(S, V, i, X, r, JX);
        if (st == 2) V.errAcc[0] = 1.0;
        if (st != 0) { S.backup[i] = 0.0; continue; }
        const double* c = S.lConst + i * S.cpf;
        const double det = c[2] * c[5] - c[3] * c[4];
        const double e0 = (c[5] * r[0] - c[3] * r[1]) / det, e1 = (-c[4] * r[0] + c[2] * r[1]) / det;
        const bool inl = e0 * e0 + e1 * e1 < thr2;
        S.backup[i] = inl ? 1.0 : 0.0;
        if (inl) acc[10] += 1.0;
        else if (skipOutliers) continue;
        const double s2 = r[0] * r[0] + r[1] * r[1];
        double soft = s2, d = 1.0;
        if (s2 > b) { const double rr = sqrt(s2); soft = 2.0 * radius * rr - b; d = radius / rr; }
        acc[0] += 0.5 * soft;
        int q = 4;
#pragma unroll
        for (int a = 0; a < 3; a++) {
          acc[1 + a] += d * (r[0] * JX[a] + r[1] * JX[3 + a]);
#pragma unroll
          for (int bb = 0; bb <= a; bb++) acc[q++] += d * (JX[a] * JX[bb] + JX[3 + a] * JX[3 + bb]);
        }
      }
    }
#pragma unroll
    for (int q = 0; q < 11; q++) acc[q] = warpSum(acc[q]);
    cost = acc[0]; nIn = acc[10];
    const double grad[3] = {acc[1], acc[2], acc[3]};
    double hess[9];
    hess[0] = acc[4]; hess[3] = hess[1] = acc[5]; hess[4] = acc[6]; hess[6] = hess[2] = acc[7]; hess[7] = hess[5] = acc[8]; hess[8] = acc[9];
    double step[3];
    if (!solve3v(hess, grad, step)) { status = -1; break; }
#pragma unroll
    for (int a = 0; a < 3; a++) X[a] -= step[a];
  }
  if (lane == 0) {
    Xg[0] = X[0]; Xg[1] = X[1]; Xg[2] = X[2];
    perPoint[3 * j] = nIn; perPoint[3 * j + 1] = cost; perPoint[3 * j + 2] = (double)status;
    if (status < 0) red_add(statsAcc, 1.0);
  }
}

void launchRefineTriangulation(const RefineArgs& A, const DevVars& v, double outlierThreshold, bool skipOutliers, int maxIts, double radius,
                               double* perPoint, double* statsAcc, cudaStream_t s) {
  if (A.nLm <= 0) return;
  refineTriangulationKernel<<<(unsigned)((A.nLm + 3) / 4), 128, 0, s>>>(A, v, outlierThreshold * outlierThreshold, skipOutliers ? 1 : 0, maxIts, radius,
                                                                         perPoint, statsAcc);
  VB_CUDA(cudaGetLastError());
}

// ---- cost only: computeSingleCost (Factor.h:390-417) + CostStats for the visual stores, over the packed landmark-major
// tables (coalesced constants, one index level).  acc = {cost, numTotal, numInvalid, numPrevInvalid}
template <bool RS>
__global__ void __launch_bounds__(256) visualCostKernel(DevStore st, DevVars V, int comparable, double* accOut) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  double acc[4] = {0, 0, 0, 0};
  if (i < st.nActive) {
    const int32_t* lv = st.pk.lVar + i * 6;
    const double* cst = st.pk.lConst + i * st.cpf;
    const int f = st.orderLm[i];
    const double prev = st.cache[f];
    double cl[8], r[2];
#pragma unroll
    for (int j = 0; j < (RS ? 8 : 6); j++) cl[j] = cst[j];
    int status;
    if (RS) status = rs_visual_factor<false>(V.data + lv[0], V.data + lv[1], V.data + lv[2], V.data + lv[3], (uint32_t)lv[5], V.data + lv[4], cl, st.aux, r,
                                             nullptr, nullptr, nullptr, nullptr, nullptr);
    else status = visual_factor<false>(V.data + lv[0], V.data + lv[1], V.data + lv[2], V.data + lv[3], cl, r, nullptr, nullptr, nullptr, nullptr) ? 0 : 1;
    if (status == 2) V.errAcc[0] = 1.0;
    const bool valid = status == 0, prevInvalid = prev < 0.0;
    acc[1] = 1.0;
    acc[2] = valid ? 0.0 : 1.0;
    acc[3] = prevInvalid ? 1.0 : 0.0;
    bool has = true;
    if (comparable) {
      if (prevInvalid) has = false;
      else if (!valid) { has = false; acc[0] = prev; }
    }
    if (has && valid) {
      double soft, d;
      loss_jet2(st.loss_a, st.loss_k, r[0] * r[0] + r[1] * r[1], soft, d);
      acc[0] = 0.5 * soft;
    }
  }
  blockReduceAddV<4>(acc, accOut);
}

void launchVisualCost(const DevStore& st, const DevVars& v, int comparable, double* acc, cudaStream_t s) {
  if (st.nActive == 0) return;
  const int grid = (int)((st.nActive + 255) / 256);
  if (st.type == VB_F_VISUAL_RS) visualCostKernel<true><<<grid, 256, 0, s>>>(st, v, comparable, acc);
  else visualCostKernel<false><<<grid, 256, 0, s>>>(st, v, comparable, acc);
  VB_CUDA(cudaGetLastError());
}

// ---- launcher ---------------------------------------------------------------------------------------------------
void launchVisualGradHess(const DevStore& st, const DevVars& v, double* grad, double* hess, int updateCache, int dontRetry,
                          double* costAcc, cudaStream_t s) {
  if (st.nActive == 0 || st.nChunks == 0) return;
  const bool rs = st.type == VB_F_VISUAL_RS;
  const size_t smemA = (size_t)(rs ? 40 : 24) * (2 * kVisChunk2 + 4) * sizeof(double);
  const size_t smemB = (size_t)4 * (2 * (6 + (rs ? 17 : 15)) + 6) * 33 * sizeof(double);
  const int gridB = (int)((st.nActive + 127) / 128);
  static bool attr = false;
  if (!attr) {
    VB_CUDA(cudaFuncSetAttribute(visualCameraKernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(24 * 260 * 8)));
    VB_CUDA(cudaFuncSetAttribute(visualCameraKernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(40 * 260 * 8)));
    VB_CUDA(cudaFuncSetAttribute(visualLandmarkKernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(4 * 48 * 33 * 8)));
    VB_CUDA(cudaFuncSetAttribute(visualLandmarkKernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(4 * 52 * 33 * 8)));
    attr = true;
  }
  if (rs) {
    visualCameraKernel<true><<<st.nChunks, 128, smemA, s>>>(st, v, grad, hess, updateCache, dontRetry, costAcc);
    visualLandmarkKernel<true><<<gridB, 128, smemB, s>>>(st, v, grad, hess, dontRetry);
  } else {
    visualCameraKernel<false><<<st.nChunks, 128, smemA, s>>>(st, v, grad, hess, updateCache, dontRetry, costAcc);
    visualLandmarkKernel<false><<<gridB, 128, smemB, s>>>(st, v, grad, hess, dontRetry);
  }
  VB_CUDA(cudaGetLastError());
}

} // namespace vb
