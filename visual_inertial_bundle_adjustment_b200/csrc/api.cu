// C ABI of libviba_b200.so (include/viba_b200.h).  Host control in C++17, all arithmetic of the LM
// iteration on the device; there is no CPU fallback: every compute entry point needs an sm_100 GPU.
#include <cstring>
#include <cstdio>
#include <cmath>
#include <algorithm>
#include <memory>
#include <climits>
#include <chrono>
#include <string>
#include "vb_internal.hpp"

using namespace vb;

static thread_local std::string g_err;
namespace vb { void setLastError(const std::string& text) { g_err = text; } }   // for the other translation units (kernels_noalloc.cu)

struct vb_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool ownsStream = true;
};

struct vb_problem {
  vb_ctx* ctx = nullptr;
  HostGraph g;
  Symbolic S;
  std::vector<void*> allocs;
  DevVars dv{};
  std::vector<DevStore> ds;
  DevSolver sol{};
  double *hess = nullptr, *grad = nullptr, *step = nullptr, *gradNew = nullptr, *backup = nullptr, *acc = nullptr;
  double* accHost = nullptr; // pinned
  int* errFlag = nullptr;
  int64_t* blockVar = nullptr;
  int64_t* diagOff = nullptr; int32_t* diagStride = nullptr; int32_t* diagSize = nullptr;
  SchurTask* schurTasks = nullptr; int nSchurTasks = 0;
  // level schedules of the supernodal kernels: [0] the lumps of this rank's own subtrees (multi-GPU only),
  // [1] the shared top of the elimination tree (one GPU: every lump)
  struct Sched {
    int32_t* lumps = nullptr; LevelTask* tasks = nullptr;
    std::vector<int64_t> lumpOff, taskOff;
    std::vector<int64_t> taskMid;   // per level: tasks [taskOff, taskMid) feed the NEXT level (urgent), [taskMid, taskOff+1) later ones
    std::vector<int32_t> maxBelow, maxWidth;
  } sched[2];
  double* dampMask = nullptr; // multi-GPU: 1 on the blocks whose diagonal gets this rank's "+ lambda"
  double* camKeep = nullptr; // multi-GPU: 1 on the camera scalars whose value this rank contributes to a sum
  int64_t varDataLen = 0;
  double visualBytes = 0;
  // shard state (vb_problem_create_sharded): this rank evaluates `owned` factors / landmarks and the
  // collective hook sums the reduced camera system over ranks
  int rank = 0, nranks = 1;
  vb_collective_fn coll = nullptr; void* collUser = nullptr;
  bool partial[3] = {false, false, true}; // step, gradNew, grad: holds only this rank's contributions
  int64_t ownedLandmarks = 0, ownedVisual = 0, collCalls = 0; double collBytes = 0;
  bool alignedPlan = true; // every contribution to a non-shared panel is local to its owner (planOwners)
  // raw IMU-0 samples and per-rig windows of the rolling-shutter stores (vb_rs_set_imu): when present the
  // tables are recomputed on the device at the top of every LM iteration
  struct RsSource {
    int store = -1;
    int64_t nRigs = 0, nSamples = 0, halfUs = 0;
    int64_t gravOff = -1;   // data offset of the gravity variable re-read at every refresh (vb_rs_set_gravity_var), -1: keep
    int64_t *timeNs = nullptr, *midUs = nullptr, *calibOff = nullptr;
    float *gyro = nullptr, *accel = nullptr;
    int* status = nullptr;
  };
  std::vector<RsSource> rsSources;
  // raw samples of one IMU and the integration interval of every factor of an inertial store (vb_preint_set_imu):
  // when present the preintegration records are recomputed on the device at the top of every LM iteration
  struct PreintSource {
    int store = -1;
    int64_t nFactors = 0, nSamples = 0;
    int64_t *timeNs = nullptr, *startUs = nullptr, *endUs = nullptr, *calibOff = nullptr;
    uint32_t* calibFlags = nullptr;
    float *gyro = nullptr, *accel = nullptr;
    double *noise6 = nullptr, *scratch = nullptr, *cov = nullptr, *omegaAtEnd = nullptr;
    int* status = nullptr;
  };
  std::vector<PreintSource> preintSources;
  // one-level lookahead of the supernodal factorisation: potrf / trsm / urgent updates on a high-priority stream, the
  // bulk of a level's updates (targets two or more levels up) on the context's stream underneath them
  cudaStream_t hiStream = nullptr;
  std::vector<cudaEvent_t> evTrsm, evLater;
  cudaEvent_t evFork = nullptr, evJoin = nullptr;
  // iterative path (vb_set_linear_solver): PCG on the reduced camera system
  int solverType = VB_SOLVER_DIRECT, pcgMaxIterations = 40, pcgLastIterations = 0;
  double pcgDesiredResidual = 1e-10, pcgLastResidual = 0.0;
  double *pcgX = nullptr, *pcgR = nullptr, *pcgZ = nullptr, *pcgP = nullptr, *pcgAp = nullptr, *pcgJac = nullptr, *pcgDots = nullptr;
  int64_t* pcgJacOff = nullptr;
  int2* pcgMvTasks = nullptr; int pcgNumMvTasks = 0;   // (supernode, row slice) of the matrix-vector product
  double* pcgPanels = nullptr;   // Gauss-Seidel / lower-precision solve: factored copy of the camera panels
  double pcgLowPrecEpsilon = 0.0;   // the diagonal raise the last lower-precision factor needed (0: none)
  double* fwdFused = nullptr;     // the vector whose forward solve the last factorAsync carried (consumed by solveAsync)
  RefineArgs refine{};            // refinePoints (vb_refine_points): landmark-major view of the visual stores
  bool refineReady = false;
  double* refinePerPoint = nullptr;
  // per refine store: its graph store, and the factor index of every slot of its landmark-major tables (host side, for
  // vb_refine_triangulation_inliers)
  std::vector<int32_t> refineStoreId;
  std::vector<std::vector<int32_t>> refineSlotFactor;
  cudaEvent_t ev[10];

  template <typename T>
  T* alloc(size_t n) {
    void* p = nullptr;
    VB_CUDA(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)));
    allocs.push_back(p);
    return (T*)p;
  }
  template <typename T>
  T* upload(const std::vector<T>& v) {
    T* p = alloc<T>(v.size());
    if (!v.empty()) VB_CUDA(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return p;
  }
  ~vb_problem() {
    for (cudaEvent_t e : evTrsm) cudaEventDestroy(e);
    for (cudaEvent_t e : evLater) cudaEventDestroy(e);
    if (evFork) cudaEventDestroy(evFork);
    if (evJoin) cudaEventDestroy(evJoin);
    if (hiStream) cudaStreamDestroy(hiStream);
    for (void* p : allocs) cudaFree(p);
    if (accHost) cudaFreeHost(accHost);
  }
};

#define VB_TRY try {
#define VB_CATCH                                                  \
  }                                                               \
  catch (const std::bad_alloc&) { g_err = "out of host memory"; return VB_ERR_STATE; } \
  catch (const std::domain_error& ex) { g_err = ex.what(); return VB_ERR_NUMERIC; }     \
  catch (const std::invalid_argument& ex) { g_err = ex.what(); return VB_ERR_ARG; }     \
  catch (const std::exception& ex) {                              \
    g_err = ex.what();                                            \
    return (g_err.find("cuda") != std::string::npos || g_err.find("CUDA") != std::string::npos) ? VB_ERR_CUDA : VB_ERR_ARG; \
  }                                                               \
  return VB_OK;

constexpr int kAccErrSlot = 63;
// landmarks that have camera-side rows, sorted by (first, last) camera position: the packing order of
// the Schur groups and the axis the shard plan cuts
static std::vector<int32_t> landmarkOrder(const Symbolic& S) {
  std::vector<int32_t> lmOrder;
  for (int64_t p = 0; p < S.nP; p++) if (S.lmRowPtr[p + 1] > S.lmRowPtr[p]) lmOrder.push_back((int32_t)p);
  std::sort(lmOrder.begin(), lmOrder.end(), [&](int32_t x, int32_t y) {
    const int32_t fx = S.lmRowCam[S.lmRowPtr[x]], fy = S.lmRowCam[S.lmRowPtr[y]];
    const int32_t lx = S.lmRowCam[S.lmRowPtr[x + 1] - 1], ly = S.lmRowCam[S.lmRowPtr[y + 1] - 1];
    return fx != fy ? fx < fy : (lx != ly ? lx < ly : x < y);
  });
  return lmOrder;
}
// Subtree-to-rank plan of the supernodal Cholesky ("the elimination tree is split into
// independent subtrees with the top separator factored on one GPU" — here on every GPU, replicated).
// Starting from the roots, the heaviest subtree is opened (its root becomes shared, its children become
// subtrees) until no subtree carries more than 1/(3 nranks) of the flops; the subtrees are then dealt to
// the ranks largest first.  The panels are re-laid out so that the shared lumps form one contiguous tail.
static void planLumpOwners(Symbolic& S, int nranks) {
  const int64_t nL = S.nL;
  S.lumpOwner.assign(nL, -1);
  S.sharedOff = S.lmDataSize; // everything shared: the whole reduced system is one summed region
  if (nranks <= 1 || nL == 0) return;
  std::vector<double> work(nL), sub(nL);
  std::vector<std::vector<int32_t>> kids(nL);
  double total = 0;
  for (int64_t L = 0; L < nL; L++) {
    const double w = S.lumpWidth[L], r = S.lumpRows[L] - S.lumpWidth[L];
    work[L] = sub[L] = w * w * w / 3.0 + w * w * r + w * r * r + 1.0;
    total += work[L];
  }
  for (int64_t L = 0; L < nL; L++) { // parents come after their children in elimination order
    const int32_t p = S.lumpParent[L];
    if (p >= 0) { if (p <= L) throw std::runtime_error("elimination tree is not topologically ordered"); sub[p] += sub[L]; kids[p].push_back((int32_t)L); }
  }
  std::vector<int32_t> roots;
  std::vector<char> shared(nL, 0), closed(nL, 0);
  for (int64_t L = 0; L < nL; L++) if (S.lumpParent[L] < 0) roots.push_back((int32_t)L);
  const double cap = total / (3.0 * nranks);
  while (true) {
    int best = -1;
    for (size_t i = 0; i < roots.size(); i++)
      if (!closed[roots[i]] && sub[roots[i]] > cap && (best < 0 || sub[roots[i]] > sub[roots[best]])) best = (int)i;
    if (best < 0) break;
    const int32_t q = roots[best];
    if (kids[q].empty()) { closed[q] = 1; continue; }
    shared[q] = 1;
    roots[best] = roots.back(); roots.pop_back();
    for (int32_t k : kids[q]) roots.push_back(k);
  }
  if ((int)roots.size() < nranks) return; // too small a tree to give every rank a subtree: replicated factorisation
  std::sort(roots.begin(), roots.end(), [&](int32_t a, int32_t b) { return sub[a] != sub[b] ? sub[a] > sub[b] : a < b; });
  std::vector<double> load(nranks, 0.0);
  std::vector<int32_t> rootOwner(nL, -1);
  for (int32_t q : roots) {
    const int r = (int)(std::min_element(load.begin(), load.end()) - load.begin());
    load[r] += sub[q];
    rootOwner[q] = r;
  }
  for (int64_t L = nL - 1; L >= 0; L--) { // owners flow down from the subtree roots
    if (shared[L]) continue;
    if (rootOwner[L] >= 0) S.lumpOwner[L] = rootOwner[L];
    else S.lumpOwner[L] = S.lumpOwner[S.lumpParent[L]];
  }
  int64_t off = (S.lmDataSize + 1) & ~int64_t(1); // same even-sized slots as buildSymbolic
  for (int pass = 0; pass < 2; pass++) {
    if (pass == 1) S.sharedOff = off;
    for (int64_t L = 0; L < nL; L++)
      if ((S.lumpOwner[L] < 0) == (pass == 1)) { S.lumpOff[L] = off; off += ((int64_t)S.lumpRows[L] * S.lumpWidth[L] + 1) & ~int64_t(1); }
  }
  if (off != S.dataSize) throw std::runtime_error("panel re-layout changed the data size");
  if (getenv("VB_VERBOSE")) {
    double sw = 0; int64_t ns = 0;
    for (int64_t L = 0; L < nL; L++) if (S.lumpOwner[L] < 0) { sw += work[L]; ns++; }
    std::fprintf(stderr, "[viba_b200] cholesky plan: %zu subtrees over %d ranks, %lld shared lumps (%.1f%% of the flops, %.1f%% of the panels), max rank load %.1f%%\n",
                 roots.size(), nranks, (long long)ns, 100.0 * sw / total, 100.0 * (S.dataSize - S.sharedOff) / std::max<int64_t>(1, S.dataSize - S.lmDataSize),
                 100.0 * *std::max_element(load.begin(), load.end()) / total);
  }
}

// Shard plan, aligned with the subtree plan of the Cholesky (planLumpOwners, called
// first).  Nested dissection separates the Schur-filled camera graph, so the cameras any landmark (or
// any factor) touches lie on ONE root-to-leaf path of the elimination tree: all of its non-shared
// cameras belong to the same rank.  That rank owns the landmark: it evaluates ALL of its observations
// and eliminates it, and every contribution to a non-shared panel is local — only the shared top of
// the reduced system is summed over ranks.  Landmarks (factors) that touch shared cameras only are
// dealt in contiguous, weight-balanced pieces (round robin for factors).  `aligned` comes back false if
// some landmark or factor straddles two ranks (not expected; the subtree panels are then summed too).
static std::vector<int32_t> planOwners(const Symbolic& S, int nranks, bool* aligned) {
  std::vector<int32_t> owner(S.nP, 0);
  if (aligned) *aligned = true;
  if (nranks <= 1) return owner;
  const std::vector<int32_t> lmOrder = landmarkOrder(S);
  auto weight = [&](int32_t p) { const double m = (double)((S.lmOff[p + 1] - S.lmOff[p]) / 3 - 3); return m * m + 300.0 * m; };
  std::vector<int32_t> loose; // landmarks that see shared cameras only
  std::vector<double> load(nranks, 0.0);
  double looseW = 0, total = 0;
  for (int32_t p : lmOrder) {
    int32_t o = -1;
    for (int64_t k = S.lmRowPtr[p]; k < S.lmRowPtr[p + 1]; k++) {
      const int32_t lo = S.lumpOwner.empty() ? -1 : S.lumpOwner[S.camLump[S.lmRowCam[k]]];
      if (lo < 0) continue;
      if (o < 0) o = lo; else if (o != lo && aligned) *aligned = false;
    }
    const double w = weight(p);
    total += w;
    if (o >= 0) { owner[p] = o; load[o] += w; } else { loose.push_back(p); looseW += w; }
  }
  // the loose landmarks fill the ranks up to the mean load, in contiguous pieces
  std::vector<double> room(nranks);
  double roomSum = 0;
  for (int r = 0; r < nranks; r++) { room[r] = std::max(0.0, total / nranks - load[r]); roomSum += room[r]; }
  if (roomSum <= 0) { for (int r = 0; r < nranks; r++) room[r] = 1.0; roomSum = nranks; }
  int cur = 0;
  double used = 0, run = 0;
  for (int32_t p : loose) {
    const double w = weight(p);
    while (cur + 1 < nranks && run + 0.5 * w > (used + room[cur]) / roomSum * looseW) { used += room[cur]; cur++; }
    owner[p] = cur;
    run += w;
  }
  for (int64_t p = 0; p < S.nP; p++) if (S.lmRowPtr[p + 1] == S.lmRowPtr[p]) owner[p] = (int32_t)(p % nranks);
  return owner;
}
// owner rank of every factor of one store: visual stores follow their landmark; the others follow the
// rank of the non-shared cameras they touch (shared only: round robin)
static std::vector<int32_t> planFactorOwners(const HostGraph& g, const Symbolic& S, const HostGraph::Store& hs,
                                             const std::vector<int32_t>& lmOwner, int nranks, bool* aligned) {
  std::vector<int32_t> out(hs.count, 0);
  if (nranks <= 1) return out;
  const bool visual = hs.type == VB_F_VISUAL || hs.type == VB_F_VISUAL_RS;
  bool ok = true;
#pragma omp parallel for schedule(static) reduction(&& : ok)
  for (int64_t k = 0; k < hs.count; k++) {
    const int32_t* v = &hs.vars[k * hs.nvars];
    if (visual) { const int64_t ii = g.index[v[0]]; out[k] = ii < 0 ? (int32_t)(k % nranks) : lmOwner[S.pos[ii]]; continue; }
    int32_t o = -1;
    for (int i = 0; i < hs.nvars; i++) {
      const int64_t ii = g.index[v[i]];
      if (ii < 0) continue;
      const int64_t pos = S.pos[ii];
      if (pos < S.nP) { const int32_t lo = lmOwner[pos]; if (o < 0) o = lo; else if (o != lo) ok = false; continue; }
      const int32_t lo = S.lumpOwner.empty() ? -1 : S.lumpOwner[S.camLump[pos - S.nP]];
      if (lo < 0) continue;
      if (o < 0) o = lo; else if (o != lo) ok = false;
    }
    out[k] = o >= 0 ? o : (int32_t)(k % nranks);
  }
  if (!ok && aligned) *aligned = false;
  return out;
}

// ---- landmark groups of the Schur update (K9b) -------------------------------------------------------
// Landmarks sorted by their first camera block are packed greedily while the union of their row blocks stays
// <= maxRows scalars and the group holds <= 32 x maxChunks landmarks; each chunk of a group is one dense
// [union rows] x [3 members] operand.  Host-only, so the padding (executed vs algorithmic flops) can be studied
// without a device (vb_schur_plan_stats).
struct SchurPlan {
  std::vector<int64_t> grpBlkPtr{0}, grpMaskPtr{0}, uniPtr{0};
  std::vector<int32_t> grpBlkOff, grpRowBlk, grpRowSub, grpRows, grpChunks, grpBlks, uniCams;
  std::vector<uint32_t> grpTileMask;
  std::vector<SchurTask> tasks;
  int64_t nGrpLm = 0;
  double executedFlops = 0; // 2 x 64 x 64 x K4 per (task, chunk with common members)
};
static void planSchurGroups(const Symbolic& S, const std::vector<int32_t>& lmOrder, int kGrpMaxRows, int kGrpChunks, SchurPlan& sp) {
  constexpr int kGrpMax = 32;
  std::vector<int32_t> uni, merged, members;
  auto scal = [&](const std::vector<int32_t>& v) { int64_t s = 0; for (int32_t c : v) s += S.camStart[c + 1] - S.camStart[c]; return s; };
  static const int kOrderMode = getenv("VB_GRP_ROW_ORDER") ? atoi(getenv("VB_GRP_ROW_ORDER")) : 1;
  static const int kMemberMode = getenv("VB_GRP_MEMBER_ORDER") ? atoi(getenv("VB_GRP_MEMBER_ORDER")) : 0;
  static const int kTile = getenv("VB_GRP_TILE_STATS") ? atoi(getenv("VB_GRP_TILE_STATS")) : 64;  // != 64: plan statistics only
  std::vector<int32_t> uniOrd, slotOf, useLo, useHi, useCnt, perm;
  auto flush = [&]() {
    if (members.empty()) return;
    const int32_t g = (int32_t)sp.grpRows.size();
    const int32_t r = (int32_t)scal(uni);
    const size_t nblk = uni.size();
    // Row order of the union.  Every 64-row tile a landmark touches costs it a full tile of DMMA work, and in
    // elimination order a member uses only ~60 % of the rows of the tiles it touches (members see different
    // subsets of rigs; velocity rows exist only for rolling-shutter observations).  Sorting the row blocks by
    // how many members use them packs the shared rows into dense leading tiles and leaves sparse tail tiles that
    // only few members are gathered for (executed / algorithmic flops at C2: 3.5 -> 2.6).  The kernel resolves
    // the triangle of every entry from the rows' elimination positions (updGlobal).
    useLo.assign(nblk, INT32_MAX); useHi.assign(nblk, -1); useCnt.assign(nblk, 0);
    for (size_t mi = 0; mi < members.size(); mi++) {
      const int32_t p = members[mi];
      size_t b = 0;
      for (int64_t k = S.lmRowPtr[p]; k < S.lmRowPtr[p + 1]; k++) {
        while (uni[b] != S.lmRowCam[k]) b++;
        useLo[b] = std::min(useLo[b], (int32_t)mi); useHi[b] = std::max(useHi[b], (int32_t)mi); useCnt[b]++;
      }
    }
    perm.resize(nblk);
    for (size_t b = 0; b < nblk; b++) perm[b] = (int32_t)b;
    if (kOrderMode == 1) { // most-shared rows first: dense leading tiles, sparse tails that few members pay for
      std::stable_sort(perm.begin(), perm.end(), [&](int32_t a, int32_t b) {
        if (useCnt[a] != useCnt[b]) return useCnt[a] > useCnt[b];
        return useLo[a] + useHi[a] < useLo[b] + useHi[b];
      });
    } else if (kOrderMode == 2) {
      std::stable_sort(perm.begin(), perm.end(), [&](int32_t a, int32_t b) { return useLo[a] + useHi[a] < useLo[b] + useHi[b]; });
    }
    uniOrd.resize(nblk); slotOf.resize(nblk);
    for (size_t i = 0; i < nblk; i++) { uniOrd[i] = uni[perm[i]]; slotOf[perm[i]] = (int32_t)i; }
    if (kMemberMode >= 1) {
      // chunks of similar extent: members re-sorted by the last (then first) slot of the sliding part of the
      // union they use, so that a chunk's 32 landmarks end in the same tiles (the tile masks are per chunk)
      int32_t nCommon = 0;
      if (kOrderMode == 1) {
        const int32_t n = (int32_t)members.size();
        for (size_t i = 0; i < nblk; i++) { if (4 * useCnt[perm[i]] >= 3 * n) nCommon++; else break; }
      }
      std::vector<std::pair<int64_t, int32_t>> keyed(members.size());
      for (size_t mi = 0; mi < members.size(); mi++) {
        const int32_t p = members[mi];
        int32_t lo = INT32_MAX, hi = -1;
        size_t b = 0;
        for (int64_t k = S.lmRowPtr[p]; k < S.lmRowPtr[p + 1]; k++) {
          while (uni[b] != S.lmRowCam[k]) b++;
          const int32_t sl = slotOf[b];
          if (sl < nCommon) continue;
          lo = std::min(lo, sl); hi = std::max(hi, sl);
        }
        keyed[mi] = {kMemberMode == 1 ? ((int64_t)hi << 32) + lo : ((int64_t)(lo + hi) << 32) + hi, p};
      }
      std::stable_sort(keyed.begin(), keyed.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
      for (size_t mi = 0; mi < members.size(); mi++) members[mi] = keyed[mi].second;
    }
    sp.uniCams.insert(sp.uniCams.end(), uniOrd.begin(), uniOrd.end());
    sp.uniPtr.push_back((int64_t)sp.uniCams.size());
    sp.grpRows.push_back(r);
    sp.nGrpLm += (int64_t)members.size();
    const int nch = ((int)members.size() + kGrpMax - 1) / kGrpMax;
    sp.grpChunks.push_back(nch);
    sp.grpBlks.push_back((int32_t)nblk);
    for (size_t b = 0; b < nblk; b++) {
      const int32_t dc = S.camStart[uniOrd[b] + 1] - S.camStart[uniOrd[b]];
      for (int32_t s = 0; s < dc; s++) { sp.grpRowBlk.push_back((int32_t)b); sp.grpRowSub.push_back(s); }
    }
    // offsets table: row (c * nblk + b), 32 member slots each
    const size_t base = sp.grpBlkOff.size();
    sp.grpBlkOff.resize(base + (size_t)nch * nblk * kGrpMax, -1);
    for (size_t mi = 0; mi < members.size(); mi++) {
      const int32_t p = members[mi];
      const int c = (int)(mi / kGrpMax), l = (int)(mi % kGrpMax);
      size_t b = 0;
      for (int64_t k = S.lmRowPtr[p]; k < S.lmRowPtr[p + 1]; k++) { // both lists ascend
        while (uni[b] != S.lmRowCam[k]) b++;
        sp.grpBlkOff[base + ((size_t)c * nblk + slotOf[b]) * kGrpMax + l] = (int32_t)(S.lmOff[p] + (int64_t)S.lmRowOff[k] * 3);
      }
    }
    sp.grpBlkPtr.push_back(sp.grpBlkPtr.back() + (int64_t)nch * (int64_t)nblk);
    // per chunk and 64-row tile: which member slots have rows in it; a tile pair only needs, chunk by
    // chunk, the landmarks present in both tiles (the others contribute exact zeros)
    const int nt = (r + kTile - 1) / kTile;
    const size_t mbase = sp.grpTileMask.size();
    sp.grpTileMask.resize(mbase + (size_t)nch * nt, 0u);
    {
      int32_t q = 0;
      for (size_t b = 0; b < nblk; b++) {
        const int32_t dc = S.camStart[uniOrd[b] + 1] - S.camStart[uniOrd[b]];
        for (int c = 0; c < nch; c++) {
          uint32_t m = 0;
          for (int l = 0; l < kGrpMax; l++) if (sp.grpBlkOff[base + ((size_t)c * nblk + b) * kGrpMax + l] >= 0) m |= 1u << l;
          for (int32_t s = 0; s < dc; s++) sp.grpTileMask[mbase + (size_t)c * nt + (q + s) / kTile] |= m;
        }
        q += dc;
      }
    }
    sp.grpMaskPtr.push_back((int64_t)sp.grpTileMask.size());
    for (int ti = 0; ti < nt; ti++)
      for (int tj = 0; tj <= ti; tj++) {
        uint32_t any = 0;
        for (int c = 0; c < nch; c++) {
          const uint32_t m = sp.grpTileMask[mbase + (size_t)c * nt + ti] & sp.grpTileMask[mbase + (size_t)c * nt + tj];
          any |= m;
          if (m) sp.executedFlops += 2.0 * kTile * kTile * ((3 * __builtin_popcount(m) + 3) & ~3);
        }
        if (any) sp.tasks.push_back({g, ti, tj, any});
      }
    uni.clear(); members.clear();
  };
  for (int32_t p : lmOrder) {
    const int32_t* bb = S.lmRowCam.data() + S.lmRowPtr[p];
    const int32_t* ee = S.lmRowCam.data() + S.lmRowPtr[p + 1];
    merged.clear();
    std::set_union(uni.begin(), uni.end(), bb, ee, std::back_inserter(merged));
    if (!members.empty() && ((int)members.size() == kGrpMax * kGrpChunks || scal(merged) > kGrpMaxRows)) {
      flush();
      merged.assign(bb, ee);
    }
    uni.swap(merged);
    members.push_back(p);
  }
  flush();
}

static void buildDeviceProblem(vb_problem& P) {
  const HostGraph& g = P.g;
  const Symbolic& S = P.S;
  // variables
  std::vector<int64_t> vecOff(g.nv, -1);
  std::vector<int64_t> blockVar(S.nB, -1);
  for (int64_t v = 0; v < g.nv; v++)
    if (g.index[v] >= 0) { vecOff[v] = S.vecStart[S.pos[g.index[v]]]; blockVar[g.index[v]] = v; }
  P.dv.nv = g.nv;
  P.dv.kind = P.upload(g.kind); P.dv.tdim = P.upload(g.tdim); P.dv.flags = P.upload(g.flags);
  P.dv.off = P.upload(g.off); P.dv.vecOff = P.upload(vecOff); P.dv.aux = P.upload(g.aux);
  P.dv.data = P.upload(g.data);
  P.varDataLen = (int64_t)g.data.size();
  P.backup = P.alloc<double>(g.data.size());
  P.blockVar = P.upload(blockVar);
  // diagonal blocks for damping
  {
    std::vector<int64_t> off(S.nB); std::vector<int32_t> st(S.nB), sz(S.nB);
    for (int64_t b = 0; b < S.nB; b++) {
      BlockRef br;
      if (!S.block(b, b, &br)) throw std::runtime_error("diagonal block missing");
      off[b] = br.off; st[b] = br.stride; sz[b] = S.bsize[b];
    }
    P.diagOff = P.upload(off); P.diagStride = P.upload(st); P.diagSize = P.upload(sz);
  }
  bool aligned = true;
  const std::vector<int32_t> owner = planOwners(S, P.nranks, &aligned);
  const bool sharded = P.nranks > 1;
  // factor stores + scatter maps
  P.ds.resize(g.stores.size());
  std::vector<std::vector<int32_t>> refLmVar(g.stores.size());   // visual stores: landmark variable of every landmark-major slot
  std::vector<std::vector<int32_t>> refSlotFactor(g.stores.size()); // ... and the factor index of the slot
  for (size_t s = 0; s < g.stores.size(); s++) {
    const HostGraph::Store& hs = g.stores[s];
    DevStore& d = P.ds[s];
    const int nv = hs.nvars, np = nv * (nv + 1) / 2;
    d.type = hs.type; d.nvars = nv; d.cpf = hs.cpf; d.npairs = np; d.count = hs.count;
    d.loss_a = hs.loss_a; d.loss_k = hs.loss_k;
    d.vars = P.upload(hs.vars); d.consts = P.upload(hs.consts); d.aux = P.upload(hs.aux);
    std::vector<double> cache(hs.count, 0.0);
    d.cache = P.upload(cache);
    std::vector<int64_t> gOff((size_t)hs.count * nv), hOff((size_t)hs.count * np);
    std::vector<int32_t> hStride((size_t)hs.count * np);
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(+ : bad)
    for (int64_t k = 0; k < hs.count; k++) {
      const int32_t* v = &hs.vars[k * nv];
      for (int i = 0; i < nv; i++) {
        const int64_t ii = g.index[v[i]];
        gOff[k * nv + i] = ii < 0 ? -1 : S.vecStart[S.pos[ii]];
        for (int j = 0; j <= i; j++) {
          const int64_t jj = g.index[v[j]];
          const int64_t pi = k * np + i * (i + 1) / 2 + j;
          hOff[pi] = -1; hStride[pi] = 0;
          if (ii < 0 || jj < 0) continue;
          BlockRef br;
          if (!S.block(ii, jj, &br)) { bad++; continue; }
          hOff[pi] = br.off;
          hStride[pi] = br.stride | ((i != j && ii != jj && br.transposed) ? (1 << 30) : 0);
        }
      }
    }
    if (bad) throw std::runtime_error("scatter map: block missing from the structure");
    d.gOff = P.upload(gOff); d.hOff = P.upload(hOff); d.hStride = P.upload(hStride);
    d.order = nullptr; d.chunkStart = nullptr; d.nChunks = 0; d.orderLm = nullptr; d.pk = VisPack{}; d.visImpl = 0;
    // Two-pass visual schedule pays once the landmark panels no longer live in L2 (measured on B200:
    // C2, 0.7 GB of panels: 6.1 ms single pass vs 6.6 ms; C3, 3 GB: 35.5 ms vs 24.6 ms).
    // VB_VISUAL_TWO_PASS=0/1 overrides.
    d.twoPass = (double)S.lmDataSize * 8.0 > 1.5e9 ? 1 : 0;
    if (const char* e = getenv("VB_VISUAL_TWO_PASS")) d.twoPass = atoi(e) != 0;
    // shard view: the factors this rank owns (planFactorOwners)
    d.kBegin = 0; d.nActive = hs.count; d.active = nullptr;
    const std::vector<int32_t> fOwner = planFactorOwners(g, S, hs, owner, P.nranks, &aligned);
    if (sharded && hs.type != VB_F_VISUAL && hs.type != VB_F_VISUAL_RS) {
      std::vector<int32_t> mine;
      for (int64_t k = 0; k < hs.count; k++) if (fOwner[k] == P.rank) mine.push_back((int32_t)k);
      d.nActive = (int64_t)mine.size();
      d.active = P.upload(mine);
    }
    if ((hs.type == VB_F_VISUAL || hs.type == VB_F_VISUAL_RS) && hs.count > 0) {
      if (hs.count >= (1ll << 31)) throw std::runtime_error("visual store too large for 32-bit factor indices");
      std::vector<int32_t> order;
      order.reserve(hs.count / P.nranks + 16);
      for (int64_t k = 0; k < hs.count; k++) if (!sharded || fOwner[k] == P.rank) order.push_back((int32_t)k);
      const int64_t nOwn = (int64_t)order.size();
      d.kBegin = 0; d.nActive = nOwn;
      if (sharded) d.active = P.upload(order); // ascending factor index: the cost kernel's order
      P.ownedVisual += nOwn;
      auto key = [&](int32_t k, int a) { return hs.vars[(int64_t)k * nv + a]; };
      std::vector<int32_t> olmKeep;
      { // landmark-major order of the second pass
        std::vector<int32_t> olm(order);
        std::stable_sort(olm.begin(), olm.end(), [&](int32_t x, int32_t y) {
          if (key(x, 0) != key(y, 0)) return key(x, 0) < key(y, 0);
          if (key(x, 2) != key(y, 2)) return key(x, 2) < key(y, 2);
          return key(x, 3) < key(y, 3);
        });
        d.orderLm = P.upload(olm);
        olmKeep.swap(olm);
      }
      // rolling shutter: the chunk also shares the velocity variable (argument 4) — one rig, one velocity in the
      // reference's builder, but the ABI does not promise it, so it is part of the key
      const bool rsStore = hs.type == VB_F_VISUAL_RS;
      std::stable_sort(order.begin(), order.end(), [&](int32_t x, int32_t y) {
        if (key(x, 1) != key(y, 1)) return key(x, 1) < key(y, 1);
        if (key(x, 2) != key(y, 2)) return key(x, 2) < key(y, 2);
        if (key(x, 3) != key(y, 3)) return key(x, 3) < key(y, 3);
        return rsStore && key(x, 4) < key(y, 4);
      });
      std::vector<int32_t> chunkStart;
      int64_t groupStart = 0;
      for (int64_t i = 0; i <= nOwn; i++) {
        const bool boundary = i == nOwn || (i > 0 && (key(order[i], 1) != key(order[i - 1], 1) ||
                              key(order[i], 2) != key(order[i - 1], 2) || key(order[i], 3) != key(order[i - 1], 3) ||
                              (rsStore && key(order[i], 4) != key(order[i - 1], 4))));
        if (boundary || i - groupStart == 128) {
          if (i > groupStart) chunkStart.push_back((int32_t)groupStart);
          groupStart = i;
        }
      }
      chunkStart.push_back((int32_t)nOwn);
      d.nChunks = (int32_t)chunkStart.size() - 1;
      d.order = P.upload(order); d.chunkStart = P.upload(chunkStart);
      // packed tables of kernels_visual.cu (VB_VISUAL_IMPL=0 keeps the round-1 kernels)
      d.visImpl = 1;
      if (const char* e = getenv("VB_VISUAL_IMPL")) d.visImpl = atoi(e);
      if (d.visImpl && (S.dataSize >= 0xffffffffll || (int64_t)g.data.size() >= 0x7fffffffll)) d.visImpl = 0;  // 32-bit offsets
      if (d.visImpl) {
        const int cpf = hs.cpf;
        const bool rs = hs.type == VB_F_VISUAL_RS;
        auto off32 = [&](int32_t var) { return (int32_t)g.off[var]; };
        std::vector<int32_t> hdr((size_t)d.nChunks * 8), cLm(nOwn);
        std::vector<double> cConst((size_t)nOwn * cpf), lConst((size_t)nOwn * cpf);
        for (int32_t c = 0; c < d.nChunks; c++) {
          const int32_t f0 = order[chunkStart[c]];
          const int32_t* v = &hs.vars[(int64_t)f0 * nv];
          int32_t* h = &hdr[(size_t)c * 8];
          h[0] = off32(v[1]); h[1] = off32(v[2]); h[2] = off32(v[3]); h[3] = rs ? off32(v[4]) : 0;
          h[4] = (int32_t)g.flags[v[3]]; h[5] = g.tdim[v[3]]; h[6] = f0; h[7] = 0;
        }
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < nOwn; i++) {
          const int32_t f = order[i];
          cLm[i] = off32(hs.vars[(int64_t)f * nv]);
          for (int j = 0; j < cpf; j++) cConst[i * cpf + j] = hs.consts[(int64_t)f * cpf + j];
        }
        std::vector<int32_t> lVar((size_t)nOwn * 6);
        std::vector<uint32_t> lOut((size_t)nOwn * 6);
        auto o32 = [](int64_t o) { return o < 0 ? 0xffffffffu : (uint32_t)o; };
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < nOwn; i++) {
          const int32_t f = olmKeep[i];
          const int32_t* v = &hs.vars[(int64_t)f * nv];
          int32_t* lv = &lVar[i * 6];
          lv[0] = off32(v[0]); lv[1] = off32(v[1]); lv[2] = off32(v[2]); lv[3] = off32(v[3]); lv[4] = rs ? off32(v[4]) : 0;
          lv[5] = (int32_t)g.flags[v[3]];
          for (int j = 0; j < cpf; j++) lConst[i * cpf + j] = hs.consts[(int64_t)f * cpf + j];
          const int64_t* ho = &hOff[(int64_t)f * np];
          uint32_t* lo = &lOut[i * 6];
          lo[0] = o32(gOff[(int64_t)f * nv]); lo[1] = o32(ho[0]); lo[2] = o32(ho[1]); lo[3] = o32(ho[3]); lo[4] = o32(ho[6]);
          lo[5] = rs ? o32(ho[10]) : 0xffffffffu;
        }
        d.pk.chunkHdr = (const int4*)P.upload(hdr); d.pk.cLm = P.upload(cLm); d.pk.cConst = P.upload(cConst);
        d.pk.lVar = P.upload(lVar); d.pk.lConst = P.upload(lConst); d.pk.lOut = P.upload(lOut);
        refLmVar[s].resize(nOwn);
        for (int64_t i = 0; i < nOwn; i++) refLmVar[s][i] = hs.vars[(int64_t)olmKeep[i] * nv];
        refSlotFactor[s].assign(olmKeep.begin(), olmKeep.begin() + nOwn);
      }
    }
    if (hs.type == VB_F_VISUAL) P.visualBytes += 224.0 * hs.count; // SURVEY.md §8d algorithmic bytes/obs
    if (hs.type == VB_F_VISUAL_RS) P.visualBytes += 296.0 * hs.count;
  }
  // refinePoints: the landmarks that have visual factors, and their range in every visual store's landmark-major tables
  if (!sharded) {
    std::vector<int32_t> lms;
    for (auto& v : refLmVar) lms.insert(lms.end(), v.begin(), v.end());
    std::sort(lms.begin(), lms.end());
    lms.erase(std::unique(lms.begin(), lms.end()), lms.end());
    int ns = 0;
    bool ok = true;
    for (size_t s = 0; s < g.stores.size(); s++) {
      if (refLmVar[s].empty()) continue;
      if (!P.ds[s].visImpl || ns >= 4) { ok = false; break; }
      const auto& lv = refLmVar[s];
      std::vector<int64_t> ptr(lms.size() + 1, 0);
      size_t i = 0;
      for (size_t j = 0; j < lms.size(); j++) {
        ptr[j] = (int64_t)i;
        while (i < lv.size() && lv[i] == lms[j]) i++;
      }
      ptr[lms.size()] = (int64_t)i;
      if (i != lv.size()) throw std::runtime_error("refinePoints: landmark-major table is not sorted by landmark");
      RefineStore& rs = P.refine.st[ns++];
      rs.lVar = P.ds[s].pk.lVar; rs.lConst = P.ds[s].pk.lConst; rs.lmPtr = P.upload(ptr);
      rs.backup = P.alloc<double>(lv.size());
      rs.aux = P.ds[s].aux; rs.loss_a = P.ds[s].loss_a; rs.loss_k = P.ds[s].loss_k; rs.cpf = P.ds[s].cpf;
      rs.rs = g.stores[s].type == VB_F_VISUAL_RS;
      P.refineStoreId.push_back((int32_t)s); P.refineSlotFactor.push_back(refSlotFactor[s]);
    }
    if (ok && ns > 0) {
      std::vector<int64_t> lmOff(lms.size());
      for (size_t j = 0; j < lms.size(); j++) lmOff[j] = g.off[lms[j]];
      P.refine.nStores = ns; P.refine.nLm = (int64_t)lms.size(); P.refine.lmOff = P.upload(lmOff);
      P.refinePerPoint = P.alloc<double>(3 * lms.size());
      P.refineReady = true;
    }
  }
  P.alignedPlan = aligned;
  if (sharded && !aligned && getenv("VB_VERBOSE")) std::fprintf(stderr, "[viba_b200] shard plan not aligned with the elimination tree: subtree panels are summed too\n");
  // solver structure
  DevSolver& d = P.sol;
  d.nP = S.nP; d.nC = S.nC; d.nL = S.nL;
  d.lmOff = P.upload(S.lmOff); d.lmRowPtr = P.upload(S.lmRowPtr); d.lmRowCam = P.upload(S.lmRowCam); d.lmRowOff = P.upload(S.lmRowOff);
  d.camLump = P.upload(S.camLump); d.camStart = P.upload(S.camStart); d.lumpFirst = P.upload(S.lumpFirst);
  d.lumpWidth = P.upload(S.lumpWidth); d.lumpRows = P.upload(S.lumpRows); d.lumpOff = P.upload(S.lumpOff);
  d.lumpRowPtr = P.upload(S.lumpRowPtr); d.lumpRowCam = P.upload(S.lumpRowCam); d.lumpRowOff = P.upload(S.lumpRowOff);
  d.camVec0 = S.vecStart[S.nP];
  {
    // relative-index tables for every scatter source: supernodes (index L) and landmark groups
    // (index nL + g).  A source is a sorted list of camera positions (its row blocks).
    std::vector<int64_t> diagLOff(S.nL), updPtr, relPtr, colOff;
    std::vector<int32_t> stride, tgt, glob, rel;
    auto appendSource = [&](const int32_t* cams, int64_t n, bool strict) {
      updPtr.push_back((int64_t)colOff.size());
      relPtr.push_back((int64_t)rel.size());
      std::vector<int32_t> targets;
      int32_t r = 0;
      for (int64_t k = 0; k < n; k++) {
        const int32_t c = cams[k];
        const int32_t T = S.camLump[c];
        int32_t t = -1; // group unions are not in elimination order: look the target up
        if (!targets.empty() && targets.back() == T) t = (int32_t)targets.size() - 1;
        else { for (size_t z = 0; z < targets.size(); z++) if (targets[z] == T) { t = (int32_t)z; break; } }
        if (t < 0) { targets.push_back(T); t = (int32_t)targets.size() - 1; }
        const int32_t dc = S.camStart[c + 1] - S.camStart[c];
        for (int32_t s = 0; s < dc; s++) {
          colOff.push_back(S.lumpOff[T] + (S.camStart[c] + s - S.camStart[S.lumpFirst[T]]));
          stride.push_back(S.lumpWidth[T]); tgt.push_back(t); glob.push_back(S.camStart[c] + s);
        }
        r += dc;
      }
      const size_t base = rel.size();
      rel.resize(base + targets.size() * (size_t)r, -1);
      for (size_t t = 0; t < targets.size(); t++) {
        const int32_t T = targets[t];
        int32_t q = 0;
        for (int64_t k = 0; k < n; k++) {
          const int32_t c = cams[k];
          const int32_t dc = S.camStart[c + 1] - S.camStart[c];
          if (c >= S.lumpFirst[T]) {
            const int32_t row = S.rowInLump(T, c);
            if (row < 0 && strict) throw std::runtime_error("symbolic: ancestor structure does not contain a descendant row");
            // (group unions: a block pair no member landmark sees together is absent and receives exact zeros)
            if (row >= 0) for (int32_t s = 0; s < dc; s++) rel[base + t * (size_t)r + q + s] = row + s;
          }
          q += dc;
        }
      }
      return r;
    };
    int64_t dl = 0;
    for (int64_t L = 0; L < S.nL; L++) {
      diagLOff[L] = dl; dl += ((int64_t)S.lumpWidth[L] * S.lumpWidth[L] + 1) & ~int64_t(1);
      appendSource(S.lumpRowCam.data() + S.lumpRowPtr[L], S.lumpRowPtr[L + 1] - S.lumpRowPtr[L], true);
    }
    // ---- landmark groups for the Schur update: landmarks sorted by their first camera block are
    // packed greedily while the union of their row blocks stays <= kGrpMaxRows scalars and the
    // group holds <= kGrpMax landmarks; each group is one dense [union rows] x [3 G] operand
    // A group is the scatter unit: up to kGrpChunks chunks of <= 32 landmarks that share one union of
    // row blocks.  The kernel accumulates the chunks of a tile pair in registers (K loop) and scatters
    // once, so the reductions per target entry drop with the number of chunks.
    int kGrpMaxRows = 1024, kGrpChunks = 4; // measured on B200 at C2: 1 chunk 5.8 ms, 4: 4.9, 8: 5.0, 16: 5.3
    if (const char* e = getenv("VB_GRP_MAX_ROWS")) kGrpMaxRows = std::max(64, atoi(e));
    if (const char* e = getenv("VB_GRP_CHUNKS")) kGrpChunks = std::max(1, std::min(32, atoi(e)));
    if (S.lmDataSize >= (1ll << 31)) throw std::runtime_error("landmark section exceeds 32-bit offsets");
    std::vector<int32_t> lmOrder = landmarkOrder(S);
    d.lmActive = nullptr; d.nLmActive = S.nP;
    if (sharded) { // keep this rank's landmarks only
      lmOrder.erase(std::remove_if(lmOrder.begin(), lmOrder.end(), [&](int32_t p) { return owner[p] != P.rank; }), lmOrder.end());
      std::vector<int32_t> mine;
      for (int64_t p = 0; p < S.nP; p++) if (owner[p] == P.rank) mine.push_back((int32_t)p);
      d.nLmActive = (int64_t)mine.size();
      d.lmActive = P.upload(mine);
    }
    P.ownedLandmarks = d.nLmActive;
    SchurPlan sp;
    planSchurGroups(S, lmOrder, kGrpMaxRows, kGrpChunks, sp);
    for (size_t g = 0; g < sp.grpRows.size(); g++) {
      const int32_t r = appendSource(sp.uniCams.data() + sp.uniPtr[g], sp.uniPtr[g + 1] - sp.uniPtr[g], false);
      if (r != sp.grpRows[g]) throw std::runtime_error("schur plan: union row count mismatch");
    }
    std::vector<int64_t>& grpBlkPtr = sp.grpBlkPtr; std::vector<int64_t>& grpMaskPtr = sp.grpMaskPtr;
    std::vector<int32_t>&grpBlkOff = sp.grpBlkOff, &grpRowBlk = sp.grpRowBlk, &grpRowSub = sp.grpRowSub, &grpRows = sp.grpRows,
                        &grpChunks = sp.grpChunks, &grpBlks = sp.grpBlks;
    std::vector<uint32_t>& grpTileMask = sp.grpTileMask;
    std::vector<SchurTask>& schurTasks = sp.tasks;
    const int64_t nGrpLm = sp.nGrpLm;
    updPtr.push_back((int64_t)colOff.size()); relPtr.push_back((int64_t)rel.size());
    d.diagLOff = P.upload(diagLOff); d.diagL = P.alloc<double>(dl);
    d.diagInv = P.alloc<double>((size_t)(S.camStart[S.nC] + 1) * 32);
    d.solveAcc = P.alloc<double>(S.camStart[S.nC] + 1); d.solveCnt = P.alloc<int>(S.nL + 1);
    VB_CUDA(cudaMemset(d.solveAcc, 0, (S.camStart[S.nC] + 1) * sizeof(double)));
    VB_CUDA(cudaMemset(d.solveCnt, 0, (S.nL + 1) * sizeof(int)));
    d.updPtr = P.upload(updPtr); d.updColOff = P.upload(colOff); d.updStride = P.upload(stride);
    d.updTgt = P.upload(tgt); d.updGlobal = P.upload(glob); d.relPtr = P.upload(relPtr); d.rel = P.upload(rel);
    d.nG = (int64_t)grpRows.size();
    d.grpBlkPtr = P.upload(grpBlkPtr); d.grpChunks = P.upload(grpChunks);
    d.grpMaskPtr = P.upload(grpMaskPtr); d.grpTileMask = P.upload(grpTileMask);
    d.grpBlkOff = P.upload(grpBlkOff); d.grpRowBlk = P.upload(grpRowBlk); d.grpRowSub = P.upload(grpRowSub);
    d.grpRows = P.upload(grpRows); d.grpBlks = P.upload(grpBlks);
    d.grpRowBase = updPtr[S.nL];
    P.schurTasks = P.upload(schurTasks); P.nSchurTasks = (int)schurTasks.size();
    if (getenv("VB_VERBOSE")) {
      double rs = 0; for (int32_t r : grpRows) rs += r;
      std::fprintf(stderr, "[viba_b200] landmark groups %zu (avg %.1f landmarks, avg %.0f union rows), schur tasks %zu, lumps %lld, levels %zu\n",
                   grpRows.size(), nGrpLm / std::max<double>(1, grpRows.size()), rs / std::max<double>(1, grpRows.size()),
                   schurTasks.size(), (long long)S.nL, S.levels.size());
    }
  }
  // level schedules
  for (int which = 0; which < 2; which++) {
    vb_problem::Sched& sc = P.sched[which];
    std::vector<int32_t> lumps; std::vector<LevelTask> tasks;
    std::vector<LevelTask> later;
    std::vector<char> tileUrgent;
    for (size_t li = 0; li < S.levels.size(); li++) {
      const auto& lv = S.levels[li];
      sc.lumpOff.push_back((int64_t)lumps.size());
      sc.taskOff.push_back((int64_t)tasks.size());
      later.clear();
      int32_t mb = 0, mw = 0;
      for (int32_t L : lv) {
        const int32_t own = S.lumpOwner.empty() ? -1 : S.lumpOwner[L];
        if (which == 0 ? own != P.rank : own >= 0) continue;
        lumps.push_back(L);
        const int r = S.lumpRows[L] - S.lumpWidth[L];
        mb = std::max(mb, r); mw = std::max(mw, S.lumpWidth[L]);
        const int nt = (r + 63) / 64;
        // a tile task (ti, tj) writes into the supernodes that own the columns of tile tj: it is urgent when one of them
        // is factored at the very next level (everything else is only needed one level later: lookahead)
        tileUrgent.assign((size_t)nt, 0);
        {
          int32_t q = 0;
          for (int64_t k = S.lumpRowPtr[L]; k < S.lumpRowPtr[L + 1]; k++) {
            const int32_t c = S.lumpRowCam[k];
            const int32_t dc = S.camStart[c + 1] - S.camStart[c];
            if (S.lumpLevel[S.camLump[c]] <= (int32_t)li + 1)
              for (int32_t t = q / 64; t <= (q + dc - 1) / 64; t++) tileUrgent[(size_t)t] = 1;
            q += dc;
          }
        }
        for (int ti = 0; ti < nt; ti++)
          for (int tj = 0; tj <= ti; tj++) (tileUrgent[(size_t)tj] ? tasks : later).push_back({L, ti, tj});
        // fused forward solve (kernels_chol.cu, kGemvSlice = 256 panel rows per task): skipped by the kernel without a
        // right-hand side; cheap, so always with the urgent part
        for (int sl = 0; sl * 256 < r; sl++) tasks.push_back({L, -1, sl});
      }
      sc.taskMid.push_back((int64_t)tasks.size());
      tasks.insert(tasks.end(), later.begin(), later.end());
      sc.maxBelow.push_back(mb); sc.maxWidth.push_back(mw);
    }
    sc.lumpOff.push_back((int64_t)lumps.size());
    sc.taskOff.push_back((int64_t)tasks.size());
    sc.lumps = P.upload(lumps); sc.tasks = P.upload(tasks);
  }
  if (sharded) { // camera scalars this rank contributes to the sums of the distributed solves
    std::vector<double> keep(S.camStart[S.nC], 0.0);
    for (int64_t L = 0; L < S.nL; L++) {
      const int32_t own = S.lumpOwner[L];
      const double k = own == P.rank || (own < 0 && P.rank == 0) ? 1.0 : 0.0;
      for (int32_t q = S.camStart[S.lumpFirst[L]]; q < S.camStart[S.lumpFirst[L + 1]]; q++) keep[q] = k;
    }
    P.camKeep = P.upload(keep);
    std::vector<double> dm(S.nB, 1.0);
    for (int64_t b = 0; b < S.nB; b++) {
      const int64_t pos = S.pos[b];
      if (pos < S.nP) { if (owner[pos] != P.rank) dm[b] = -1.0; continue; } // not this rank's landmark: untouched
      const int32_t own = S.lumpOwner[S.camLump[pos - S.nP]];
      dm[b] = own == P.rank || (own < 0 && P.rank == 0) ? 1.0 : 0.0;
    }
    P.dampMask = P.upload(dm);
  }
  P.hess = P.alloc<double>(S.dataSize);
  P.grad = P.alloc<double>(S.order); P.step = P.alloc<double>(S.order); P.gradNew = P.alloc<double>(S.order);
  P.acc = P.alloc<double>(64 + 4 * g.stores.size());
  P.errFlag = P.alloc<int>(1);
  P.dv.errAcc = P.acc + kAccErrSlot;
  VB_CUDA(cudaMallocHost(&P.accHost, (64 + 4 * g.stores.size()) * sizeof(double)));
  VB_CUDA(cudaMemset(P.hess, 0, S.dataSize * sizeof(double)));
  VB_CUDA(cudaMemset(P.grad, 0, S.order * sizeof(double)));
  VB_CUDA(cudaMemset(P.step, 0, S.order * sizeof(double)));
  VB_CUDA(cudaMemset(P.gradNew, 0, S.order * sizeof(double)));
  VB_CUDA(cudaMemset(P.errFlag, 0, sizeof(int)));
  for (auto& e : P.ev) VB_CUDA(cudaEventCreate(&e));
}

// scalar accumulator slots in p->acc: the reducible ones are contiguous so that a sharded run sums
// them over ranks with one collective.  [16,20) applyStep ratio stats (identical on every rank).
constexpr int kAccDot = 61, kAccCost = 62, kAccErr = 63, kAccStores = 64;
// collective hook: in-place sum over ranks / broadcast from rank 0, ordered on the problem's stream
static void collective(vb_problem* p, int op, double* ptr, int64_t n) {
  if (p->nranks <= 1 || n <= 0) return;
  if (!p->coll) throw std::runtime_error("sharded problem without a collective hook");
  if (p->coll(p->collUser, op, ptr, n, (void*)p->ctx->stream) != 0) throw std::runtime_error("collective hook failed");
  p->collCalls++; p->collBytes += 8.0 * n;
}
static double* vecOf(vb_problem* p, int which) { return which == 0 ? p->step : which == 1 ? p->gradNew : p->grad; }
// every ABI entry runs on the problem's device, whatever the calling thread's current device is
static void requireProblem(const vb_problem* p) {
  if (!p) throw std::invalid_argument("null problem");
  VB_CUDA(cudaSetDevice(p->ctx->device));
}
static void syncCheck(vb_problem* p) { VB_CUDA(cudaStreamSynchronize(p->ctx->stream)); }

constexpr int64_t kRsTableStride = 1 + 32 * 11 + 31 * 9; // per-rig table (dev_factors_ext.cuh)

// RollingShutterData::compute for every rig of the stores that were given IMU samples
// (InitCalibration.cpp:316-325, called from the preStepCallback: Optimizer.cpp:801-805)
// computePreIntegration for every factor of the stores that have raw samples attached
static bool launchPreint(vb_problem* p, int store) {
  cudaStream_t s = p->ctx->stream;
  bool any = false;
  for (auto& src : p->preintSources) {
    if (store >= 0 && src.store != store) continue;
    any = true;
    VB_CUDA(cudaMemsetAsync(src.status, 0, sizeof(int), s));
    launchPreintRefresh(src.timeNs, src.gyro, src.accel, src.nSamples, p->dv.data, src.calibOff, src.calibFlags, src.startUs, src.endUs,
                        src.nFactors, src.noise6, src.scratch, const_cast<double*>(p->ds[src.store].aux), src.cov, src.omegaAtEnd, src.status, s);
    VB_CUDA(cudaGetLastError());
  }
  return any;
}
static void checkPreint(vb_problem* p, int store) {
  cudaStream_t s = p->ctx->stream;
  for (auto& src : p->preintSources) {
    if (store >= 0 && src.store != store) continue;
    int st = 0;
    VB_CUDA(cudaMemcpyAsync(&st, src.status, sizeof(int), cudaMemcpyDeviceToHost, s));
    VB_CUDA(cudaStreamSynchronize(s));
    if (st == 1) throw std::domain_error("enumIntegrationSteps: not enough IMU samples around a preintegration interval");
    if (st != 0) throw std::domain_error("computePreIntegration: empty interval");
  }
}
static bool launchRsTables(vb_problem* p, int store) {
  cudaStream_t s = p->ctx->stream;
  bool any = false;
  for (auto& src : p->rsSources) {
    if (store >= 0 && src.store != store) continue;
    any = true;
    VB_CUDA(cudaMemsetAsync(src.status, 0, sizeof(int), s));
    // gravityWorld_ of every RollingShutterData is re-snapshotted from the gravity variable (InitCalibration.cpp:317,
    // RollingShutterData.cpp:21): constant in single-session runs, a free S2 variable in multi-session ones
    if (src.gravOff >= 0)
      VB_CUDA(cudaMemcpyAsync(const_cast<double*>(p->ds[src.store].aux), p->dv.data + src.gravOff, 3 * sizeof(double), cudaMemcpyDeviceToDevice, s));
    double* tables = const_cast<double*>(p->ds[src.store].aux) + 4;
    launchRsRefresh(src.timeNs, src.gyro, src.accel, src.nSamples, p->dv.data, src.calibOff, src.midUs, src.halfUs, src.nRigs,
                    tables, src.status, s);
    VB_CUDA(cudaGetLastError());
  }
  return any;
}
// the reference throws from inside compute(); here the status word is read back when the host next waits
static void checkRsTables(vb_problem* p, int store) {
  cudaStream_t s = p->ctx->stream;
  for (auto& src : p->rsSources) {
    if (store >= 0 && src.store != store) continue;
    int st = 0;
    VB_CUDA(cudaMemcpyAsync(&st, src.status, sizeof(int), cudaMemcpyDeviceToHost, s));
    VB_CUDA(cudaStreamSynchronize(s));
    if (st == 1) throw std::domain_error("enumIntegrationSteps: not enough IMU samples around a rolling-shutter window");
    if (st == 2) throw std::domain_error("RollingShutterData::compute: more than 32 samples in a window");
    if (st != 0) throw std::domain_error("RollingShutterData::compute: sample times not increasing");
  }
}
static void refreshRsTables(vb_problem* p, int store) {
  if (!launchRsTables(p, store) && store >= 0) throw std::invalid_argument("store has no IMU samples (vb_rs_set_imu)");
  checkRsTables(p, store);
}

extern "C" {

int vb_version(void) { return 100; }
const char* vb_last_error(void) { return g_err.c_str(); }

void vb_settings_default(vb_settings* s) { // lib/small_thing/Optimizer.h:40-78
  if (!s) return;
  s->damping = 1e-5; s->damping_adjust_on_fail = 2.5; s->damping_adjust_on_good = 0.7; s->damping_adjust_on_average = 1.5;
  s->damping_min = 1e-9; s->damping_max = 1e8;
  s->min_relative_cost_reduction = 0.3; s->step_factor_decrease = 0.3; s->min_step_factor_for_good = 0.7;
  s->absolute_cost_tolerance = 1e-8; s->relative_cost_tolerance = 1e-10; s->variables_tolerance = 1e-5;
  s->max_step_factor_attempts = 2; s->try_sub_step = 1;
  s->stop_if_no_improvement_for = 3; s->distance_from_troubled_iteration = 3; s->max_num_iterations = 50;
  s->verbose = 0;
  s->solver_type = VB_SOLVER_DIRECT; s->pcg_max_iterations = 40; s->pcg_desired_residual = 1e-10;
}

int vb_ctx_create(int device, vb_ctx** out) {
  VB_TRY
  if (!out) throw std::invalid_argument("null out");
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) throw std::runtime_error("CUDA: no device available (this library has no CPU fallback)");
  if (device < 0 || device >= n) throw std::invalid_argument("device index out of range");
  cudaDeviceProp prop;
  VB_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) throw std::runtime_error("CUDA: device is not sm_100 (Blackwell) — kernels are built for sm_100a only");
  VB_CUDA(cudaSetDevice(device));
  auto* c = new vb_ctx();
  c->device = device;
  VB_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  *out = c;
  VB_CATCH
}
int vb_ctx_create_on_stream(int device, void* cuda_stream, vb_ctx** out) {
  const int rc = vb_ctx_create(device, out);
  if (rc != VB_OK) return rc;
  cudaStreamDestroy((*out)->stream);
  (*out)->stream = (cudaStream_t)cuda_stream; // caller-owned (e.g. the stream its NCCL communicator orders against)
  (*out)->ownsStream = false;
  return VB_OK;
}
int vb_ctx_destroy(vb_ctx* ctx) {
  if (!ctx) return VB_OK;
  if (ctx->ownsStream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return VB_OK;
}

int vb_problem_create(vb_ctx* ctx, const vb_graph_desc* g, vb_problem** out) {
  return vb_problem_create_sharded(ctx, g, 0, 1, nullptr, nullptr, out);
}
int vb_problem_create_sharded(vb_ctx* ctx, const vb_graph_desc* g, int rank, int nranks, vb_collective_fn fn, void* user, vb_problem** out) {
  VB_TRY
  if (!ctx || !g || !out) throw std::invalid_argument("null argument");
  if (nranks < 1 || rank < 0 || rank >= nranks) throw std::invalid_argument("rank / nranks out of range");
  if (nranks > 1 && !fn) throw std::invalid_argument("a sharded problem needs a collective hook");
  VB_CUDA(cudaSetDevice(ctx->device));
  std::unique_ptr<vb_problem> P(new vb_problem());
  P->ctx = ctx;
  P->rank = rank; P->nranks = nranks; P->coll = fn; P->collUser = user;
  P->partial[2] = nranks > 1;
  P->g.load(g);
  SymbolicOptions opt;
  if (const char* e = getenv("VB_MAX_LUMP_WIDTH")) opt.maxLumpWidth = std::max(8, std::min(128, atoi(e)));
  if (const char* e = getenv("VB_ND_LEAF_BLOCKS")) opt.ndLeafBlocks = atoi(e);
  buildSymbolic(P->g, opt, &P->S);
  planLumpOwners(P->S, getenv("VB_REPLICATED_CHOLESKY") ? 1 : nranks);
  buildDeviceProblem(*P);
  *out = P.release();
  VB_CATCH
}
int vb_problem_destroy(vb_problem* p) { delete p; return VB_OK; }

// host-only shard plan: owner rank of every landmark (registration index) for an nranks-way split
static void hostPlan(const vb_graph_desc* g, int nranks, HostGraph& hg, Symbolic& S) {
  hg.load(g);
  SymbolicOptions opt;
  if (const char* e = getenv("VB_MAX_LUMP_WIDTH")) opt.maxLumpWidth = std::max(8, std::min(128, atoi(e)));
  if (const char* e = getenv("VB_ND_LEAF_BLOCKS")) opt.ndLeafBlocks = atoi(e);
  buildSymbolic(hg, opt, &S);
  planLumpOwners(S, getenv("VB_REPLICATED_CHOLESKY") ? 1 : nranks);
}
int vb_shard_plan(const vb_graph_desc* g, int nranks, int32_t* landmark_owner, int64_t n) {
  VB_TRY
  if (!g || !landmark_owner || nranks < 1) throw std::invalid_argument("bad argument");
  HostGraph hg; Symbolic S;
  hostPlan(g, nranks, hg, S);
  if (n != S.nP) throw std::invalid_argument("length must be num_points");
  const std::vector<int32_t> owner = planOwners(S, nranks, nullptr);
  for (int64_t i = 0; i < S.nP; i++) landmark_owner[i] = owner[S.pos[i]];
  VB_CATCH
}
int vb_shard_factor_owners(const vb_graph_desc* g, int nranks, int32_t store, int32_t* factor_owner, int64_t n, int32_t* aligned) {
  VB_TRY
  if (!g || !factor_owner || nranks < 1 || store < 0 || store >= g->num_stores) throw std::invalid_argument("bad argument");
  HostGraph hg; Symbolic S;
  hostPlan(g, nranks, hg, S);
  if (n != hg.stores[store].count) throw std::invalid_argument("length must be the store's factor count");
  bool al = true;
  const std::vector<int32_t> owner = planOwners(S, nranks, &al);
  const std::vector<int32_t> fo = planFactorOwners(hg, S, hg.stores[store], owner, nranks, &al);
  std::copy(fo.begin(), fo.end(), factor_owner);
  if (aligned) *aligned = al ? 1 : 0;
  VB_CATCH
}
// {rank, nranks, owned landmarks, owned visual factors, doubles of the reduced camera system summed
//  over ranks per factorisation, collective calls so far, collective bytes so far, lumps in this rank's own subtrees}
int vb_shard_info(const vb_problem* p, double* out8) {
  VB_TRY
  requireProblem(p);
  out8[0] = p->rank; out8[1] = p->nranks; out8[2] = (double)p->ownedLandmarks; out8[3] = (double)p->ownedVisual;
  out8[4] = (double)(p->S.dataSize - p->S.lmDataSize); out8[5] = (double)p->collCalls; out8[6] = p->collBytes;
  out8[7] = (double)p->sched[0].lumpOff.back(); // lumps of this rank's own subtrees (0 on one GPU)
  VB_CATCH
}

// symbolic-only entry (host): lets CPU-side tests check the structure without a GPU
int vb_symbolic_only(const vb_graph_desc* g, int64_t* order, int64_t* dataSize, int64_t* nLumps, int64_t* perm, int64_t permLen) {
  VB_TRY
  HostGraph hg; hg.load(g);
  Symbolic S; buildSymbolic(hg, SymbolicOptions(), &S);
  if (order) *order = S.order;
  if (dataSize) *dataSize = S.dataSize;
  if (nLumps) *nLumps = S.nL;
  if (perm) { if (permLen != S.nB) throw std::invalid_argument("perm length"); std::copy(S.pos.begin(), S.pos.end(), perm); }
  VB_CATCH
}

// host-only: {order, dataSize, lumps, levels, maxLumpRows, cholFlops, schurFlops, widest level (lumps)}
int vb_symbolic_stats(const vb_graph_desc* g, int max_lump_width, int nd_leaf_blocks, double* out8) {
  VB_TRY
  HostGraph hg; hg.load(g);
  SymbolicOptions opt;
  if (max_lump_width > 0) opt.maxLumpWidth = max_lump_width;
  opt.ndLeafBlocks = nd_leaf_blocks;
  Symbolic S; buildSymbolic(hg, opt, &S);
  size_t widest = 0;
  for (const auto& lv : S.levels) widest = std::max(widest, lv.size());
  out8[0] = (double)S.order; out8[1] = (double)S.dataSize; out8[2] = (double)S.nL; out8[3] = (double)S.levels.size();
  out8[4] = S.maxLumpRows; out8[5] = S.cholFlops; out8[6] = S.schurFlops; out8[7] = (double)widest;
  VB_CATCH
}

// host-only: padding of the Schur update for a grouping (max union rows, chunks of 32 landmarks per group):
// out8 = {groups, tasks, executed DMMA flops, algorithmic Schur flops, mean landmarks per group, mean union rows,
//         levels, algorithmic Cholesky flops}
int vb_schur_plan_stats(const vb_graph_desc* g, int max_rows, int chunks, int nd_leaf_blocks, double* out8) {
  VB_TRY
  HostGraph hg; hg.load(g);
  SymbolicOptions opt;
  if (nd_leaf_blocks != 0) opt.ndLeafBlocks = nd_leaf_blocks;
  if (const char* e = getenv("VB_MAX_LUMP_WIDTH")) opt.maxLumpWidth = std::max(8, std::min(128, atoi(e)));
  Symbolic S; buildSymbolic(hg, opt, &S);
  SchurPlan sp;
  planSchurGroups(S, landmarkOrder(S), max_rows, chunks, sp);
  double rs = 0; for (int32_t r : sp.grpRows) rs += r;
  const double ng = std::max<double>(1, (double)sp.grpRows.size());
  out8[0] = (double)sp.grpRows.size(); out8[1] = (double)sp.tasks.size(); out8[2] = sp.executedFlops; out8[3] = S.schurFlops;
  out8[4] = sp.nGrpLm / ng; out8[5] = rs / ng; out8[6] = (double)S.levels.size(); out8[7] = S.cholFlops;
  VB_CATCH
}

int64_t vb_order(const vb_problem* p) { return p ? p->S.order : -1; }
int64_t vb_data_size(const vb_problem* p) { return p ? p->S.dataSize : -1; }
int64_t vb_num_blocks(const vb_problem* p) { return p ? p->S.nB : -1; }
int64_t vb_num_lumps(const vb_problem* p) { return p ? p->S.nL : -1; }
int64_t vb_num_levels(const vb_problem* p) { return p ? (int64_t)p->S.levels.size() : -1; }
int64_t vb_num_launches_per_iteration(const vb_problem* p) { // kernels one vb_lm_iteration_timed launches
  if (!p) return -1;
  int64_t n = 0;
  for (const auto& st : p->ds) n += st.count > 0 ? (st.orderLm && st.twoPass ? 3 : 2) : 0; // grad/Hess (visual: two passes) + cost
  int64_t lv = 0;
  for (const auto& sc : p->sched) for (size_t i = 0; i + 1 < sc.lumpOff.size(); i++) lv += sc.lumpOff[i + 1] > sc.lumpOff[i];
  n += (int64_t)p->rsSources.size(); // rolling-shutter table refresh
  n += (int64_t)p->preintSources.size(); // preintegration refresh
  return n + 2 /*damp, step=grad*/ + 1 + 2 * lv /*factor*/ + 2 + 2 * lv /*solve*/ + 2 /*dot, negate*/ + 1 /*apply*/;
}
int vb_get_param_start(const vb_problem* p, int64_t* out, int64_t n) {
  VB_TRY
  requireProblem(p);
  if (n != p->S.nB) throw std::invalid_argument("length must be vb_num_blocks");
  for (int64_t i = 0; i < n; i++) out[i] = p->S.vecStart[p->S.pos[i]];
  VB_CATCH
}
int vb_get_perm(const vb_problem* p, int64_t* out, int64_t n) {
  VB_TRY
  requireProblem(p);
  if (n != p->S.nB) throw std::invalid_argument("length must be vb_num_blocks");
  std::copy(p->S.pos.begin(), p->S.pos.end(), out);
  VB_CATCH
}
int vb_get_lump_start(const vb_problem* p, int64_t* out, int64_t n) {
  VB_TRY
  requireProblem(p);
  if (n != p->S.nL + 1) throw std::invalid_argument("length must be vb_num_lumps + 1");
  for (int64_t i = 0; i < n; i++) out[i] = p->S.nP + p->S.lumpFirst[i];
  VB_CATCH
}
int vb_block_offset(const vb_problem* p, int64_t i, int64_t j, int64_t* offset, int64_t* stride, int32_t* transposed) {
  VB_TRY
  requireProblem(p);
  if (i < 0 || j < 0 || i >= p->S.nB || j >= p->S.nB) throw std::invalid_argument("block index out of range");
  BlockRef br;
  if (!p->S.block(i, j, &br)) { *offset = -1; *stride = 0; *transposed = 0; }
  else { *offset = br.off; *stride = br.stride; *transposed = br.transposed; }
  VB_CATCH
}
int vb_get_work(const vb_problem* p, double* schur_flops, double* chol_flops, double* visual_bytes) {
  VB_TRY
  requireProblem(p);
  if (schur_flops) *schur_flops = p->S.schurFlops;
  if (chol_flops) *chol_flops = p->S.cholFlops;
  if (visual_bytes) *visual_bytes = p->visualBytes;
  VB_CATCH
}

static double evalGradHess(vb_problem* p, double* gvec, int with_hess, int update_cache, int dont_retry, bool sync) {
  cudaStream_t s = p->ctx->stream;
  VB_CUDA(cudaMemsetAsync(gvec, 0, p->S.order * sizeof(double), s));
  if (with_hess) {
    if (p->nranks > 1) { // only what this rank writes: its landmarks' panels and the camera part
      launchZeroOwnedPanels(p->sol, p->hess, s);
      VB_CUDA(cudaMemsetAsync(p->hess + p->S.lmDataSize, 0, (p->S.dataSize - p->S.lmDataSize) * sizeof(double), s));
    } else {
      VB_CUDA(cudaMemsetAsync(p->hess, 0, p->S.dataSize * sizeof(double), s));
    }
  }
  VB_CUDA(cudaMemsetAsync(p->acc + kAccCost, 0, 2 * sizeof(double), s)); // cost + evaluation error word
  for (const DevStore& st : p->ds) launchGradHess(st, p->dv, gvec, with_hess ? p->hess : nullptr, update_cache, dont_retry, p->acc + kAccCost, s);
  p->partial[gvec == p->step ? 0 : gvec == p->gradNew ? 1 : 2] = p->nranks > 1;
  if (!sync) return 0.0;
  collective(p, VB_COLL_SUM, p->acc + kAccCost, 2);
  VB_CUDA(cudaMemcpyAsync(p->accHost + kAccCost, p->acc + kAccCost, 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
  syncCheck(p);
  if (p->accHost[kAccErr] != 0.0) throw std::domain_error("RollingShutterData::getEstimate: time shift out of the sampled range");
  return p->accHost[kAccCost];
}
int vb_eval_grad_hess(vb_problem* p, int with_hess, int update_cache, int dont_retry_failed, double* cost) {
  VB_TRY
  requireProblem(p);
  const double c = evalGradHess(p, p->grad, with_hess, update_cache, dont_retry_failed, true);
  if (cost) *cost = c;
  VB_CATCH
}
int vb_eval_grad_new(vb_problem* p, int dont_retry_failed) {
  VB_TRY
  requireProblem(p);
  evalGradHess(p, p->gradNew, 0, 0, dont_retry_failed, true);
  VB_CATCH
}
static double evalCost(vb_problem* p, int comparable, vb_cost_stats* stats) {
  cudaStream_t s = p->ctx->stream;
  const size_t n = p->ds.size();
  double* acc = p->acc + kAccStores;
  VB_CUDA(cudaMemsetAsync(p->acc + kAccErr, 0, (1 + 4 * n) * sizeof(double), s));
  for (size_t i = 0; i < n; i++) launchCost(p->ds[i], p->dv, comparable, acc + 4 * i, s);
  collective(p, VB_COLL_SUM, p->acc + kAccErr, 1 + 4 * (int64_t)n);
  VB_CUDA(cudaMemcpyAsync(p->accHost + kAccErr, p->acc + kAccErr, (1 + 4 * n) * sizeof(double), cudaMemcpyDeviceToHost, s));
  syncCheck(p);
  if (p->accHost[kAccErr] != 0.0) throw std::domain_error("RollingShutterData::getEstimate: time shift out of the sampled range");
  double tot = 0;
  for (size_t i = 0; i < n; i++) tot += p->accHost[kAccStores + 4 * i];
  if (stats && n) { // Q3: CostStats of the last store only (Factor.h:686-688, Optimizer.cpp:93-95)
    const double* a = p->accHost + kAccStores + 4 * (n - 1);
    stats->num_total = (int64_t)a[1]; stats->num_invalid = (int64_t)a[2]; stats->num_prev_invalid = (int64_t)a[3];
  }
  return tot;
}
int vb_eval_cost(vb_problem* p, int comparable, double* cost, vb_cost_stats* stats) {
  VB_TRY
  requireProblem(p);
  const double c = evalCost(p, comparable, stats);
  if (cost) *cost = c;
  VB_CATCH
}
int vb_damp(vb_problem* p, double lambda) {
  VB_TRY
  requireProblem(p);
  launchDamping(p->hess, p->diagOff, p->diagStride, p->diagSize, p->S.nB, lambda, p->dampMask, p->ctx->stream);
  VB_CATCH
}
int vb_step_from_grad(vb_problem* p) {
  VB_TRY
  requireProblem(p);
  launchVecOp(0, p->step, p->grad, 0, p->S.order, nullptr, p->ctx->stream);
  p->partial[0] = p->partial[2];
  VB_CATCH
}

static void factorLevels(vb_problem* p, int which, double* rhs = nullptr) {
  cudaStream_t s = p->ctx->stream;
  const vb_problem::Sched& sc = p->sched[which];
  const size_t nLv = sc.lumpOff.size() - 1;
  static const bool kLookahead = !(getenv("VB_LOOKAHEAD") && atoi(getenv("VB_LOOKAHEAD")) == 0);
  if (!kLookahead) {
    for (size_t lv = 0; lv < nLv; lv++) {
      const int nl = (int)(sc.lumpOff[lv + 1] - sc.lumpOff[lv]);
      const int nt = (int)(sc.taskOff[lv + 1] - sc.taskOff[lv]);
      if (nl == 0) continue;
      launchLumpFactor(p->sol, p->hess, sc.lumps + sc.lumpOff[lv], nl, sc.maxBelow[lv], sc.maxWidth[lv], p->errFlag, s, rhs);
      launchLumpUpdate(p->sol, p->hess, sc.tasks + sc.taskOff[lv], nt, sc.maxWidth[lv], nl > 1, s, rhs);
    }
    return;
  }
  // One-level lookahead.  The potrf -> trsm -> urgent-update chain of level k runs on a high-priority stream; the rest of
  // level k's update tasks (targets at level k + 2 and above) run on the context's stream underneath the chain of level
  // k + 1, which only waits for the "later" tasks of level k - 1.  Targets may now be hit by two levels at once: always
  // fp64 reductions.
  if (!p->hiStream) {
    int least = 0, greatest = 0;
    VB_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    VB_CUDA(cudaStreamCreateWithPriority(&p->hiStream, cudaStreamNonBlocking, greatest));
    VB_CUDA(cudaEventCreateWithFlags(&p->evFork, cudaEventDisableTiming));
    VB_CUDA(cudaEventCreateWithFlags(&p->evJoin, cudaEventDisableTiming));
  }
  while (p->evTrsm.size() < nLv) {
    cudaEvent_t a, b;
    VB_CUDA(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
    VB_CUDA(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
    p->evTrsm.push_back(a); p->evLater.push_back(b);
  }
  cudaStream_t h = p->hiStream;
  VB_CUDA(cudaEventRecord(p->evFork, s));
  VB_CUDA(cudaStreamWaitEvent(h, p->evFork, 0));
  std::vector<long> laterLevels;   // levels whose "later" tasks were queued on s, ascending
  size_t waited = 0;               // laterLevels[0 .. waited) have been waited for by h
  for (size_t lv = 0; lv < nLv; lv++) {
    const int nl = (int)(sc.lumpOff[lv + 1] - sc.lumpOff[lv]);
    if (nl == 0) continue;
    const int nUrgent = (int)(sc.taskMid[lv] - sc.taskOff[lv]);
    const int nLater = (int)(sc.taskOff[lv + 1] - sc.taskMid[lv]);
    // this level's supernodes need every update from two or more levels below: s is in order, so waiting for the most
    // recent such batch of later tasks covers the older ones
    size_t upTo = waited;
    while (upTo < laterLevels.size() && laterLevels[upTo] + 2 <= (long)lv) upTo++;
    if (upTo > waited) { VB_CUDA(cudaStreamWaitEvent(h, p->evLater[(size_t)laterLevels[upTo - 1]], 0)); waited = upTo; }
    launchLumpFactor(p->sol, p->hess, sc.lumps + sc.lumpOff[lv], nl, sc.maxBelow[lv], sc.maxWidth[lv], p->errFlag, h, rhs);
    if (nLater > 0) VB_CUDA(cudaEventRecord(p->evTrsm[lv], h));
    launchLumpUpdate(p->sol, p->hess, sc.tasks + sc.taskOff[lv], nUrgent, sc.maxWidth[lv], 1, h, rhs);
    if (nLater > 0) {
      VB_CUDA(cudaStreamWaitEvent(s, p->evTrsm[lv], 0));
      launchLumpUpdate(p->sol, p->hess, sc.tasks + sc.taskMid[lv], nLater, sc.maxWidth[lv], 1, s, nullptr);
      VB_CUDA(cudaEventRecord(p->evLater[lv], s));
      laterLevels.push_back((long)lv);
    }
  }
  VB_CUDA(cudaEventRecord(p->evJoin, h));
  VB_CUDA(cudaStreamWaitEvent(s, p->evJoin, 0));
}
// rhs (one GPU only): the forward half of Solver::solve on that vector is fused into the factorisation — the landmark
// forward solve runs between the landmark Cholesky and the Schur update, every supernode's segment rides through its trsm
// and its panel product through the update kernel — so the 2 x levels dependent kernels of a separate forward pass
// disappear from the iteration's critical path.  solveAsync then only runs the backward half (p->fwdFused).
static void factorPcgAsync(vb_problem* p, cudaEvent_t afterSchur);
static void solvePcg(vb_problem* p, double* vec);
static void factorAsync(vb_problem* p, cudaEvent_t afterSchur, double* rhs = nullptr) {
  if (p->solverType != VB_SOLVER_DIRECT) { factorPcgAsync(p, afterSchur); return; }
  cudaStream_t s = p->ctx->stream;
  const Symbolic& S = p->S;
  VB_CUDA(cudaMemsetAsync(p->errFlag, 0, sizeof(int), s));
  static const bool kFuse = !(getenv("VB_FUSE_FWD") && atoi(getenv("VB_FUSE_FWD")) == 0);
  if ((p->nranks > 1 && !p->alignedPlan) || !kFuse) rhs = nullptr;
  p->fwdFused = rhs;
  if (rhs) {
    launchSchurPrep(p->sol, p->hess, p->errFlag, s);
    launchSolveLandmarksFwd(p->sol, p->hess, rhs, s);
    launchSchurUpdate(p->sol, p->hess, p->schurTasks, p->nSchurTasks, s);
    if (afterSchur) VB_CUDA(cudaEventRecord(afterSchur, s));
    if (p->nranks > 1) {
      // Sharded, tree-aligned plan.  The right-hand side is rank-partial: an own-subtree camera entry is complete on its
      // owner and exactly zero elsewhere (no factor and no landmark of another rank touches it), a shared-top entry holds
      // this rank's part.  After the own subtrees are factored (forward solve fused) ONE sum of the camera vector gives
      // every rank the solved subtrees of all ranks and the complete right-hand side of the shared top.
      factorLevels(p, 0, rhs);
      collective(p, VB_COLL_SUM, p->hess + S.sharedOff, S.dataSize - S.sharedOff);
      collective(p, VB_COLL_SUM, rhs + p->sol.camVec0, S.order - p->sol.camVec0);
    }
    factorLevels(p, 1, rhs);
    return;
  }
  launchSchur(p->sol, p->hess, p->schurTasks, p->nSchurTasks, p->errFlag, s);
  // sharded: every rank eliminated its own landmarks from its own partial camera system; the sum over
  // ranks is the reduced camera system (the exchange step of SURVEY.md §8e).  With the tree-aligned
  // shard plan every contribution to the panels of a rank-owned subtree is already local to that rank;
  // the shared top keeps accumulating partial sums (local Schur part + the updates of this rank's
  // subtrees) and is summed once, after the subtrees are factored.
  if (!p->alignedPlan) collective(p, VB_COLL_SUM, p->hess + S.lmDataSize, S.sharedOff - S.lmDataSize);
  if (afterSchur) VB_CUDA(cudaEventRecord(afterSchur, s));
  if (p->nranks > 1) {
    factorLevels(p, 0);
    collective(p, VB_COLL_SUM, p->hess + S.sharedOff, S.dataSize - S.sharedOff);
  }
  factorLevels(p, 1);
}
static void checkNumeric(vb_problem* p) {
  int flag = 0;
  VB_CUDA(cudaMemcpyAsync(&flag, p->errFlag, sizeof(int), cudaMemcpyDeviceToHost, p->ctx->stream));
  syncCheck(p);
  if (p->nranks > 1) { // every rank must leave the iteration together: share the verdict
    p->accHost[60] = (double)flag;
    VB_CUDA(cudaMemcpyAsync(p->acc + 60, p->accHost + 60, sizeof(double), cudaMemcpyHostToDevice, p->ctx->stream));
    collective(p, VB_COLL_SUM, p->acc + 60, 1);
    VB_CUDA(cudaMemcpyAsync(p->accHost + 60, p->acc + 60, sizeof(double), cudaMemcpyDeviceToHost, p->ctx->stream));
    syncCheck(p);
    if (flag == 0 && p->accHost[60] != 0.0) throw std::domain_error("factor: not positive definite on another rank");
  }
  if (flag == 1) throw std::domain_error("factor: landmark block not positive definite");
  if (flag == 2) throw std::domain_error("factor: reduced camera system not positive definite");
}
static void solveLevels(vb_problem* p, double* vec, int which, bool forward) {
  cudaStream_t s = p->ctx->stream;
  const vb_problem::Sched& sc = p->sched[which];
  const int nLv = (int)sc.lumpOff.size() - 1;
  for (int i = 0; i < nLv; i++) {
    const int lv = forward ? i : nLv - 1 - i;
    const int nl = (int)(sc.lumpOff[lv + 1] - sc.lumpOff[lv]);
    if (nl == 0) continue;
    if (forward) launchSolveLumpsFwd(p->sol, p->hess, vec, sc.lumps + sc.lumpOff[lv], nl, sc.maxWidth[lv], sc.maxBelow[lv], nl > 1, s);
    else launchSolveLumpsBwd(p->sol, p->hess, vec, sc.lumps + sc.lumpOff[lv], nl, sc.maxWidth[lv], sc.maxBelow[lv], s);
  }
}
static void solveAsync(vb_problem* p, double* vec) {
  if (p->solverType != VB_SOLVER_DIRECT) { solvePcg(p, vec); return; }
  cudaStream_t s = p->ctx->stream;
  const int64_t nCam = p->S.order - p->sol.camVec0;
  double* cam = vec + p->sol.camVec0;
  if (p->fwdFused == vec) {   // the forward half already happened inside factorAsync
    p->fwdFused = nullptr;
    solveLevels(p, vec, 1, false);
    if (p->nranks > 1) {
      solveLevels(p, vec, 0, false);
      launchVecOp(4, cam, p->camKeep, 0, nCam, nullptr, s);   // every camera scalar from exactly one rank
      collective(p, VB_COLL_SUM, cam, nCam);
    }
    launchSolveLandmarksBwd(p->sol, p->hess, vec, s);
    collective(p, VB_COLL_SUM, vec, p->sol.camVec0);          // owned landmark steps -> full step everywhere
    p->partial[vec == p->step ? 0 : vec == p->gradNew ? 1 : 2] = false;
    return;
  }
  launchSolveLandmarksFwd(p->sol, p->hess, vec, s);
  if (p->nranks > 1) {
    collective(p, VB_COLL_SUM, cam, nCam);           // partial camera gradients - owned landmark eliminations
    // forward over this rank's subtrees: keep only the scalars this rank contributes (its subtrees; the
    // shared top on rank 0), so that one sum yields the solved subtrees of every rank and the complete
    // right-hand side of the shared top
    launchVecOp(4, cam, p->camKeep, 0, nCam, nullptr, s);
    solveLevels(p, vec, 0, true);
    collective(p, VB_COLL_SUM, cam, nCam);
  }
  solveLevels(p, vec, 1, true);
  solveLevels(p, vec, 1, false);
  if (p->nranks > 1) {
    solveLevels(p, vec, 0, false);
    // every camera scalar from exactly one rank: one bit pattern of the camera step everywhere
    launchVecOp(4, cam, p->camKeep, 0, nCam, nullptr, s);
    collective(p, VB_COLL_SUM, cam, nCam);
  }
  launchSolveLandmarksBwd(p->sol, p->hess, vec, s);
  collective(p, VB_COLL_SUM, vec, p->sol.camVec0);   // owned landmark steps -> full step everywhere
  p->partial[vec == p->step ? 0 : vec == p->gradNew ? 1 : 2] = false;
}
// ---- iterative path (Optimizer.cpp:232-331) ---------------------------------------------------------------------------------
static void pcgEnsureBuffers(vb_problem* p) {
  if (p->pcgX) return;
  const Symbolic& S = p->S;
  const int64_t nCam = S.order - p->sol.camVec0;
  p->pcgX = p->alloc<double>(nCam); p->pcgR = p->alloc<double>(nCam); p->pcgZ = p->alloc<double>(nCam);
  p->pcgP = p->alloc<double>(nCam); p->pcgAp = p->alloc<double>(nCam);
  p->pcgDots = p->alloc<double>(8);
  std::vector<int64_t> off((size_t)S.nC + 1, 0);
  for (int64_t c = 0; c < S.nC; c++) { const int64_t d = S.camStart[c + 1] - S.camStart[c]; off[(size_t)c + 1] = off[(size_t)c] + d * d; }
  p->pcgJacOff = p->upload(off);
  p->pcgJac = p->alloc<double>(off[(size_t)S.nC]);
  std::vector<int2> tasks;
  for (int64_t L = 0; L < S.nL; L++)
    for (int sl = 0; sl * camSymMvRows() < S.lumpRows[L]; sl++) tasks.push_back(make_int2((int)L, sl));
  p->pcgMvTasks = p->upload(tasks); p->pcgNumMvTasks = (int)tasks.size();
}
// factorUpTo(elimination range) + precond->init (Optimizer.cpp:289-300)
static void factorPcgAsync(vb_problem* p, cudaEvent_t afterSchur) {
  if (p->nranks > 1) throw std::logic_error("the PCG path is not available on a sharded problem");
  cudaStream_t s = p->ctx->stream;
  pcgEnsureBuffers(p);
  p->fwdFused = nullptr;
  VB_CUDA(cudaMemsetAsync(p->errFlag, 0, sizeof(int), s));
  launchSchur(p->sol, p->hess, p->schurTasks, p->nSchurTasks, p->errFlag, s);
  if (afterSchur) VB_CUDA(cudaEventRecord(afterSchur, s));
  if (p->solverType == VB_SOLVER_PCG_JACOBI) launchJacobiInit(p->sol, p->hess, p->pcgJacOff, p->pcgJac, p->errFlag, s);
  if (p->solverType != VB_SOLVER_PCG_GAUSS_SEIDEL && p->solverType != VB_SOLVER_PCG_LOWER_PREC_SOLVE) return;
  // Both preconditioners work on a copy of the unfactored camera panels, addressed through a "virtual" base pointer
  // exactly as the reference does it (Preconditioner.h:131-137,183-186): the supernode kernels only touch panel
  // offsets >= lmDataSize.
  const Symbolic& S = p->S;
  const int64_t camDoubles = S.dataSize - S.lmDataSize;
  if (!p->pcgPanels) p->pcgPanels = p->alloc<double>(camDoubles);
  double* virt = p->pcgPanels - S.lmDataSize;
  const vb_problem::Sched& sc = p->sched[1];
  VB_CUDA(cudaMemcpyAsync(p->pcgPanels, p->hess + S.lmDataSize, camDoubles * sizeof(double), cudaMemcpyDeviceToDevice, s));
  if (p->solverType == VB_SOLVER_PCG_GAUSS_SEIDEL) {
    // BlockGaussSeidelPrecond::init = Solver::pseudoFactorFrom: the Cholesky sweep without its Schur updates — every
    // supernode's diagonal block factored, its panel multiplied by the inverse transpose — so that the two supernodal
    // solves apply ((D + E) D^-1 (D + E)^T)^-1, D = the supernode diagonal blocks.  No dependencies: one launch pair.
    const int nl = (int)(sc.lumpOff.back() - sc.lumpOff.front());
    int maxBelow = 0, maxWidth = 1;
    for (size_t lv = 0; lv + 1 < sc.lumpOff.size(); lv++) { maxBelow = std::max(maxBelow, (int)sc.maxBelow[lv]); maxWidth = std::max(maxWidth, (int)sc.maxWidth[lv]); }
    if (nl > 0) launchLumpFactor(p->sol, virt, sc.lumps + sc.lumpOff.front(), nl, maxBelow, maxWidth, p->errFlag, s, nullptr);
    return;
  }
  // LowerPrecSolvePrecond::init (Preconditioner.h:178-218): complete factor of the float copy; while it fails the
  // diagonal is raised, d -> d (1 + eps) + eps with eps = 1e-8, 3e-8, ...  Here the DATA is rounded to float and the
  // factorisation runs on the fp64 tensor path; "fails" = a non-positive pivot instead of the reference's NaN test.
  double epsilon = 0.0;
  p->pcgLowPrecEpsilon = 0.0;
  int landmarkFlag = 0;
  VB_CUDA(cudaMemcpyAsync(&landmarkFlag, p->errFlag, sizeof(int), cudaMemcpyDeviceToHost, s));
  syncCheck(p);
  if (landmarkFlag != 0) return;   // the landmark elimination failed: checkNumeric reports it
  for (int attempt = 0; attempt < 64; attempt++) {
    if (attempt > 0) VB_CUDA(cudaMemcpyAsync(p->pcgPanels, p->hess + S.lmDataSize, camDoubles * sizeof(double), cudaMemcpyDeviceToDevice, s));
    launchRoundToFloat(p->pcgPanels, camDoubles, s);
    if (epsilon > 0) { launchCamDiagRaise(p->sol, virt, epsilon, s); p->pcgLowPrecEpsilon = epsilon; epsilon *= 3.0; }
    else epsilon = 1e-8;
    VB_CUDA(cudaMemsetAsync(p->errFlag, 0, sizeof(int), s));
    std::swap(p->hess, virt);
    try { factorLevels(p, 1); } catch (...) { std::swap(p->hess, virt); throw; }
    std::swap(p->hess, virt);
    int flag = 0;
    VB_CUDA(cudaMemcpyAsync(&flag, p->errFlag, sizeof(int), cudaMemcpyDeviceToHost, s));
    syncCheck(p);
    if (flag == 0) return;
  }
  // still failing: errFlag stays set and checkNumeric turns it into VB_ERR_NUMERIC / a rejected step
}
// solveLUpTo + PCG::solve on the reduced system + solveLtUpTo (Optimizer.cpp:301-327, PCG.cpp:15-97)
static void solvePcg(vb_problem* p, double* vec) {
  cudaStream_t s = p->ctx->stream;
  const Symbolic& S = p->S;
  const int64_t nCam = S.order - p->sol.camVec0;
  double* cam = vec + p->sol.camVec0;
  double *x = p->pcgX, *r = p->pcgR, *z = p->pcgZ, *pk = p->pcgP, *Ap = p->pcgAp, *dots = p->pcgDots;
  const bool jacobi = p->solverType == VB_SOLVER_PCG_JACOBI;
  const bool factored = p->solverType == VB_SOLVER_PCG_GAUSS_SEIDEL || p->solverType == VB_SOLVER_PCG_LOWER_PREC_SOLVE;
  const bool lowPrec = p->solverType == VB_SOLVER_PCG_LOWER_PREC_SOLVE;
  auto precond = [&](const double* in, double* out) {
    if (jacobi) { launchJacobiApply(p->sol, p->pcgJacOff, p->pcgJac, in, out, s); return; }
    VB_CUDA(cudaMemcpyAsync(out, in, nCam * sizeof(double), cudaMemcpyDeviceToDevice, s));
    if (!factored) return;
    // solveLFrom + solveLtFrom on the (pseudo-)factored copy (Preconditioner.h:139-153, 220-239); the supernode solve
    // kernels address vec + camVec0 + ..., so the nCam-long vector goes in behind a virtual base as well
    double* virt = p->pcgPanels - S.lmDataSize;
    if (lowPrec) launchRoundToFloat(out, nCam, s);
    std::swap(p->hess, virt);
    try {
      solveLevels(p, out - p->sol.camVec0, 1, true);
      solveLevels(p, out - p->sol.camVec0, 1, false);
    } catch (...) { std::swap(p->hess, virt); throw; }
    std::swap(p->hess, virt);
    if (lowPrec) launchRoundToFloat(out, nCam, s);
  };
  auto readDots = [&](double* host, int n) {
    VB_CUDA(cudaMemcpyAsync(host, dots, n * sizeof(double), cudaMemcpyDeviceToHost, s));
    syncCheck(p);
  };
  launchSolveLandmarksFwd(p->sol, p->hess, vec, s);
  VB_CUDA(cudaMemsetAsync(x, 0, nCam * sizeof(double), s));
  VB_CUDA(cudaMemcpyAsync(r, cam, nCam * sizeof(double), cudaMemcpyDeviceToDevice, s));
  precond(r, z);
  VB_CUDA(cudaMemcpyAsync(pk, z, nCam * sizeof(double), cudaMemcpyDeviceToDevice, s));
  // dots: [0] z.r (current), [1] r.r, [2] p.Ap, [3] scratch, [4] z1.r1 (next)
  VB_CUDA(cudaMemsetAsync(dots, 0, 8 * sizeof(double), s));
  launchPcgDot2(z, r, r, r, nCam, dots, s);
  double h[8];
  readDots(h, 2);
  const double rk0norm = std::sqrt(h[1]);
  double rel = 0.0;
  int k = 0;
  if (!(rk0norm > 0.0)) { p->pcgLastIterations = 0; p->pcgLastResidual = 0.0; }
  else while (true) {
    VB_CUDA(cudaMemsetAsync(Ap, 0, nCam * sizeof(double), s));
    launchCamSymMv(p->sol, p->hess, p->pcgMvTasks, p->pcgNumMvTasks, pk, Ap, S.maxLumpWidth, s);
    VB_CUDA(cudaMemsetAsync(dots + 2, 0, 2 * sizeof(double), s));
    launchPcgDot2(pk, Ap, pk, Ap, nCam, dots + 2, s);
    launchPcgStep(x, r, pk, Ap, dots + 0, dots + 2, nCam, s);       // alpha = z.r / p.Ap, on the device
    precond(r, z);                                                   // z_{k+1} (one more than the reference when it stops here)
    VB_CUDA(cudaMemsetAsync(dots + 4, 0, 2 * sizeof(double), s));
    launchPcgDot2(z, r, r, r, nCam, dots + 4, s);
    readDots(h, 6);
    rel = std::sqrt(h[5]) / rk0norm;
    if (!std::isfinite(rel)) break;   // the reference's PCG does not check either; the LM loop rejects the useless step
    if (rel < p->pcgDesiredResidual) break;
    if (k + 1 >= p->pcgMaxIterations) break;
    launchPcgDir(pk, z, dots + 4, dots + 0, nCam, s);                // beta = z1.r1 / z.r
    VB_CUDA(cudaMemcpyAsync(dots, dots + 4, sizeof(double), cudaMemcpyDeviceToDevice, s));
    k++;
  }
  if (rk0norm > 0.0) { p->pcgLastIterations = k + 1; p->pcgLastResidual = rel; }
  VB_CUDA(cudaMemcpyAsync(cam, x, nCam * sizeof(double), cudaMemcpyDeviceToDevice, s));
  launchSolveLandmarksBwd(p->sol, p->hess, vec, s);
  p->partial[vec == p->step ? 0 : vec == p->gradNew ? 1 : 2] = false;
}
#define VB_CATCH_NUMERIC                                                             \
  }                                                                                  \
  catch (const std::domain_error& ex) { g_err = ex.what(); return VB_ERR_NUMERIC; } \
  catch (const std::invalid_argument& ex) { g_err = ex.what(); return VB_ERR_ARG; }  \
  catch (const std::exception& ex) { g_err = ex.what(); return VB_ERR_CUDA; }        \
  return VB_OK;

static const char* solverName(int solverType) {   // Optimizer::solverToString (Optimizer.cpp:148-164)
  switch (solverType) {
    case VB_SOLVER_DIRECT: return "direct";
    case VB_SOLVER_PCG_TRIVIAL: return "trivial";
    case VB_SOLVER_PCG_JACOBI: return "jacobi";
    case VB_SOLVER_PCG_GAUSS_SEIDEL: return "gauss-seidel";
    case VB_SOLVER_PCG_LOWER_PREC_SOLVE: return "lower-prec-solve";
    default: return "<unknown>";
  }
}
static void setLinearSolver(vb_problem* p, int solverType, int maxIterations, double desiredResidual) {
  if (solverType < VB_SOLVER_DIRECT || solverType > VB_SOLVER_PCG_LOWER_PREC_SOLVE)
    throw std::invalid_argument("Unknown preconditioner " + std::to_string(solverType));   // Optimizer.cpp:243-245
  if (solverType != VB_SOLVER_DIRECT) {
    if (p->nranks > 1) throw std::invalid_argument("the PCG path runs on an unsharded problem only");
    if (maxIterations < 1 || !(desiredResidual >= 0.0)) throw std::invalid_argument("pcg_max_iterations >= 1 and pcg_desired_residual >= 0 required");
    p->pcgMaxIterations = maxIterations; p->pcgDesiredResidual = desiredResidual;
  }
  if (solverType != p->solverType) p->fwdFused = nullptr;
  p->solverType = solverType;
}
int vb_set_linear_solver(vb_problem* p, int32_t solver_type, int32_t pcg_max_iterations, double pcg_desired_residual) {
  VB_TRY
  requireProblem(p);
  setLinearSolver(p, solver_type, pcg_max_iterations, pcg_desired_residual);
  VB_CATCH_NUMERIC
}
int vb_pcg_last(const vb_problem* p, int32_t* iterations, double* residual) {
  VB_TRY
  requireProblem(const_cast<vb_problem*>(p));
  if (iterations) *iterations = p->pcgLastIterations;
  if (residual) *residual = p->pcgLastResidual;
  VB_CATCH_NUMERIC
}
int vb_factor(vb_problem* p) {
  VB_TRY
  requireProblem(p);
  factorAsync(p, nullptr);
  checkNumeric(p);
  VB_CATCH_NUMERIC
}
int vb_solve(vb_problem* p, int which) {
  VB_TRY
  requireProblem(p);
  solveAsync(p, vecOf(p, which));
  syncCheck(p);
  VB_CATCH_NUMERIC
}
static bool isPartial(const vb_problem* p, const double* v) { return p->partial[v == p->step ? 0 : v == p->gradNew ? 1 : 2]; }
// a.b; sharded: a full (replicated) vector against a partial one is the sum of the local dots
static double dotVec(vb_problem* p, const double* a, const double* b) {
  cudaStream_t s = p->ctx->stream;
  const bool pa = isPartial(p, a), pb = isPartial(p, b);
  if (pa && pb) throw std::logic_error("dot of two rank-partial vectors is not defined on a sharded problem");
  VB_CUDA(cudaMemsetAsync(p->acc + kAccDot, 0, sizeof(double), s));
  launchVecOp(2, const_cast<double*>(a), b, 0, p->S.order, p->acc + kAccDot, s);
  if (pa != pb) collective(p, VB_COLL_SUM, p->acc + kAccDot, 1);
  VB_CUDA(cudaMemcpyAsync(p->accHost + kAccDot, p->acc + kAccDot, sizeof(double), cudaMemcpyDeviceToHost, s));
  syncCheck(p);
  return p->accHost[kAccDot];
}
int vb_step_stats(vb_problem* p, double* step_dot_grad, double* grad_norm, double* step_norm) {
  VB_TRY
  requireProblem(p);
  if (step_dot_grad) *step_dot_grad = dotVec(p, p->step, p->grad);
  if (grad_norm) { // needs the full gradient: sum the rank-partial vectors first (idempotent afterwards)
    if (p->partial[2]) { collective(p, VB_COLL_SUM, p->grad, p->S.order); p->partial[2] = false; }
    *grad_norm = std::sqrt(dotVec(p, p->grad, p->grad));
  }
  if (step_norm) *step_norm = std::sqrt(dotVec(p, p->step, p->step));
  VB_CATCH
}
int vb_scale_vec(vb_problem* p, int which, double sc) {
  VB_TRY
  requireProblem(p);
  launchVecOp(1, vecOf(p, which), nullptr, sc, p->S.order, nullptr, p->ctx->stream);
  VB_CATCH
}
int vb_dot(vb_problem* p, int a, int b, double* out) {
  VB_TRY
  requireProblem(p);
  if (!out) throw std::invalid_argument("null out");
  *out = dotVec(p, vecOf(p, a), vecOf(p, b));
  VB_CATCH
}
static void applyStep(vb_problem* p, const double* vec, double ratios[3]) {
  cudaStream_t s = p->ctx->stream;
  VB_CUDA(cudaMemsetAsync(p->acc + 16, 0, 4 * sizeof(double), s));
  launchApplyStep(p->dv, p->blockVar, p->S.nB, vec, p->acc + 16, s);
  if (!ratios) return;
  VB_CUDA(cudaMemcpyAsync(p->accHost + 16, p->acc + 16, 4 * sizeof(double), cudaMemcpyDeviceToHost, s));
  syncCheck(p);
  const double n = p->accHost[19];
  ratios[0] = p->accHost[16]; ratios[1] = std::sqrt(p->accHost[17] / n); ratios[2] = p->accHost[18] / n;
}
int vb_apply_step(vb_problem* p, int which, double ratios[3]) {
  VB_TRY
  requireProblem(p);
  double r[3];
  applyStep(p, vecOf(p, which), r);
  if (ratios) { ratios[0] = r[0]; ratios[1] = r[1]; ratios[2] = r[2]; }
  VB_CATCH
}
int vb_backup(vb_problem* p) {
  VB_TRY
  requireProblem(p);
  VB_CUDA(cudaMemcpyAsync(p->backup, p->dv.data, p->varDataLen * sizeof(double), cudaMemcpyDeviceToDevice, p->ctx->stream));
  VB_CATCH
}
int vb_restore(vb_problem* p) {
  VB_TRY
  requireProblem(p);
  VB_CUDA(cudaMemcpyAsync(p->dv.data, p->backup, p->varDataLen * sizeof(double), cudaMemcpyDeviceToDevice, p->ctx->stream));
  VB_CATCH
}

// text helpers of the per-iteration log (the layout of lib/small_thing/Print.cpp: a duration as
// "123μs" / "1.23ms" / "0.12s" / "1m5s" / "1h2m", a ratio as a percentage with fixed precision)
static std::string fmtDuration(double seconds) {
  char buf[64];
  const long long us = (long long)std::llround(seconds * 1e6);
  if (us < 1000) std::snprintf(buf, sizeof buf, "%lld\xce\xbcs", us);
  else if (us < 100000) std::snprintf(buf, sizeof buf, "%.2fms", us / 1e3);
  else if (us < 60000000ll) std::snprintf(buf, sizeof buf, "%.2fs", us / 1e6);
  else if (us < 3600000000ll) {
    const long long m = us / 60000000ll, sec = std::llround((us % 60000000ll) / 1e6);
    if (sec > 0) std::snprintf(buf, sizeof buf, "%lldm%llds", m, sec); else std::snprintf(buf, sizeof buf, "%lldm", m);
  } else {
    const long long h = us / 3600000000ll, m = (us % 3600000000ll) / 60000000ll;
    if (m > 0) std::snprintf(buf, sizeof buf, "%lldh%lldm", h, m); else std::snprintf(buf, sizeof buf, "%lldh", h);
  }
  return buf;
}
static std::string fmtPercent(double ratio, int precision = 1) {
  char buf[64];
  std::snprintf(buf, sizeof buf, "%.*f%%", precision, ratio * 100.0);
  return buf;
}
static std::string fmtG(double v) { // operator<< of a double on a default stream: %g with 6 significant digits
  char buf[64];
  std::snprintf(buf, sizeof buf, "%g", v);
  return buf;
}

// Optimizer::optimize — lib/small_thing/Optimizer.cpp:768-1106, host control over device state.
// Quirks Q1-Q5 of SURVEY.md App. A are kept (see the inline notes).  Hooks: Settings::preStepCallback
// (Optimizer.h:90, Optimizer.cpp:803-805), Settings::log (Optimizer.h:81) and a per-iteration trace.
int vb_optimize_ex(vb_problem* p, const vb_settings* st, vb_optimize_hooks* hooks, vb_summary* out) {
  double initialCost = 0, finalCost = 0;
  int numTroubledSeqs = 0, largestTroubledSeq = 0, stopReason = 1, iterationNum = 0;
  auto fillSummary = [&]() {
    if (!out) return;
    out->initial_cost = initialCost; out->final_cost = finalCost;
    out->num_troubled_seqs = numTroubledSeqs; out->largest_troubled_seq = largestTroubledSeq;
    out->num_iterations = iterationNum; out->stop_reason = stopReason;
  };
  try {
  requireProblem(p);
  if (!st) throw std::invalid_argument("null settings");
  if (hooks && hooks->pre_step && hooks->sync_kinds && !hooks->host_vars) throw std::invalid_argument("sync_kinds without host_vars");
  if (hooks) hooks->trace_len = 0;
  setLinearSolver(p, st->solver_type, st->pcg_max_iterations, st->pcg_desired_residual);   // Optimizer.cpp:774 (initSolverData)
  using clk = std::chrono::steady_clock;
  auto since = [](clk::time_point a) { return std::chrono::duration<double>(clk::now() - a).count(); };
  cudaStream_t s = p->ctx->stream;
  const int64_t n = p->S.order;
  const bool wantLog = hooks && hooks->log;
  const bool wantTrace = hooks && hooks->trace && hooks->trace_cap > 0;
  auto say = [&](const std::string& text) { if (wantLog) hooks->log(hooks->log_user, text.c_str()); };
  double damping = st->damping;
  int lastImprovementIteration = 0, lastTroubledIteration = -10;
  double troubledSeqStartDamping = damping;
  int troubledSeqStart = 0;
  bool dontRetryFailed = false;
  auto acceptable = [](const vb_cost_stats& c) {
    const double rate = c.num_invalid / (c.num_total + 1.0);
    return rate < 0.03 && (c.num_invalid < c.num_prev_invalid * 2.0 + 50);
  };
  while (true) {
    const auto startIt = clk::now();
    int branch = 0, attempts = 0;
    if (hooks && hooks->pre_step) { // Optimizer.cpp:803-805; the host copy of the kinds the callback reads comes first
      if (hooks->sync_kinds) {
        syncCheck(p);
        const HostGraph& g = p->g;
        int64_t runBeg = -1, runEnd = -1; // contiguous runs of selected variables -> one copy each
        auto flush = [&]() {
          if (runBeg >= 0) VB_CUDA(cudaMemcpyAsync(hooks->host_vars + runBeg, p->dv.data + runBeg, (runEnd - runBeg) * sizeof(double), cudaMemcpyDeviceToHost, s));
          runBeg = -1;
        };
        for (int64_t v = 0; v < g.nv; v++) {
          if (!((hooks->sync_kinds >> g.kind[v]) & 1u)) continue;
          const int64_t b = g.off[v], e = v + 1 < g.nv ? g.off[v + 1] : p->varDataLen;
          if (b != runEnd) { flush(); runBeg = b; }
          runEnd = e;
        }
        flush();
        syncCheck(p);
      }
      hooks->pre_step(hooks->pre_step_user, iterationNum);
    }
    // the reference's own preStepCallback (main_AriaKit_ViBa.cpp:95-101): preintegrations first, then the RS tables
    if (!p->preintSources.empty()) { launchPreint(p, -1); checkPreint(p, -1); }
    if (!p->rsSources.empty()) refreshRsTables(p, -1); // InitCalibration.cpp:316-325
    const double prevCost = evalGradHess(p, p->grad, 1, 1, dontRetryFailed, true);
    const double secGradHess = since(startIt); // includes the callback, as Optimizer.cpp:801,816 measures it
    finalCost = prevCost;
    if (iterationNum == 0) initialCost = prevCost;
    const double dampingUsed = damping;
    launchDamping(p->hess, p->diagOff, p->diagStride, p->diagSize, p->S.nB, damping, p->dampMask, s);
    launchVecOp(0, p->step, p->grad, 0, n, nullptr, s);
    const auto startFactor = clk::now();
    factorAsync(p, nullptr, p->step);
    try {
      checkNumeric(p);
    } catch (const std::domain_error& ex) {
      // A pivot of the damped system was not positive in floating point.  The reference's loop never aborts: a useless
      // step is rejected and the damping raised (Optimizer.cpp:1019-1035).  Same here — the variables were not touched
      // yet — instead of returning VB_ERR_NUMERIC mid-run.
      p->fwdFused = nullptr;
      if (lastTroubledIteration != iterationNum - 1) { troubledSeqStartDamping = damping; troubledSeqStart = iterationNum; }
      say(std::string("factorisation failed (") + ex.what() + "): step rejected, damping raised");
      damping *= st->damping_adjust_on_fail;
      if (wantTrace && hooks->trace_len < hooks->trace_cap) {
        vb_iter_trace& t = hooks->trace[hooks->trace_len++];
        std::memset(&t, 0, sizeof t);
        t.prev_cost = prevCost; t.new_cost = prevCost; t.damping = damping; t.damping_used = dampingUsed;
        t.branch = VB_BR_REJECTED | VB_BR_FACTOR_FAILED;
      }
      if (damping > st->damping_max) { say("damping out of range, quadratic model failing?!"); stopReason = 2; break; }
      lastTroubledIteration = iterationNum;
      iterationNum++;
      continue;
    }
    const double secFactor = since(startFactor);
    const auto startSolve = clk::now();
    solveAsync(p, p->step);
    const double modelCostReduction = dotVec(p, p->step, p->grad) * 0.5;
    const double secSolve = since(startSolve);
    if (modelCostReduction < 0) { // Q1: recompute, bump damping, fall through with the bad step
      evalGradHess(p, p->grad, 1, 1, dontRetryFailed, true);
      damping *= st->damping_adjust_on_fail;
      branch |= VB_BR_MODEL_NEGATIVE;
      say(" ?:# quadratic model failing numerically, retrying... (damping: " + fmtG(damping) + ")");
    }
    launchVecOp(1, p->step, nullptr, -1.0, n, nullptr, s);
    double gradNorm = 0, stepNorm = 0;
    if (wantLog || wantTrace) { // Optimizer.cpp:858-859 (only the log reads them)
      if (p->partial[2]) { collective(p, VB_COLL_SUM, p->grad, n); p->partial[2] = false; }
      gradNorm = std::sqrt(dotVec(p, p->grad, p->grad));
      stepNorm = std::sqrt(dotVec(p, p->step, p->step));
    }
    VB_CUDA(cudaMemcpyAsync(p->backup, p->dv.data, p->varDataLen * sizeof(double), cudaMemcpyDeviceToDevice, s));
    double ratios[3];
    applyStep(p, p->step, ratios);
    vb_cost_stats stats{};
    double newCost = evalCost(p, 1, &stats);
    double costReduction = prevCost - newCost;
    const double ratioReductionToCost = costReduction / newCost;
    double ratioReductionToExpected = costReduction / modelCostReduction;
    double appliedStepFactor = 1.0;
    bool failureRateIsAcceptable = acceptable(stats);
    if (st->max_step_factor_attempts > 0 &&
        (ratioReductionToExpected < st->min_relative_cost_reduction || !failureRateIsAcceptable)) {
      branch |= VB_BR_STEP_FACTOR;
      evalGradHess(p, p->gradNew, 0, 0, dontRetryFailed, false);
      const double backRed = -dotVec(p, p->gradNew, p->step) * 0.5;
      double stepFactor = backRed > 0 ? (modelCostReduction / (modelCostReduction + backRed)) : st->step_factor_decrease;
      for (int i = 0; i < st->max_step_factor_attempts; i++) {
        attempts++;
        appliedStepFactor *= stepFactor;
        launchVecOp(1, p->step, nullptr, stepFactor, n, nullptr, s);
        VB_CUDA(cudaMemcpyAsync(p->dv.data, p->backup, p->varDataLen * sizeof(double), cudaMemcpyDeviceToDevice, s));
        applyStep(p, p->step, nullptr);
        vb_cost_stats statsAtF{};
        const double newCostAtF = evalCost(p, 1, &statsAtF);
        const double costReductionAtF = prevCost - newCost; // Q2: stale newCost, as in the reference
        const double ratioAtF = costReductionAtF / (modelCostReduction * appliedStepFactor);
        char fbuf[32];
        std::snprintf(fbuf, sizeof fbuf, "%.2f", appliedStepFactor);
        if (ratioAtF >= st->min_relative_cost_reduction && acceptable(statsAtF)) {
          newCost = newCostAtF; stats = statsAtF; costReduction = costReductionAtF;
          ratioReductionToExpected = ratioAtF; failureRateIsAcceptable = true;
          branch |= VB_BR_ACCEPT_AT_FACTOR;
          say(std::string("\\!/ cost reduction obtained applying factor ") + fbuf);
          break;
        }
        if (st->try_sub_step) {
          branch |= VB_BR_SUB_STEP;
          evalGradHess(p, p->gradNew, 0, 0, dontRetryFailed, false);
          const auto startSub = clk::now();
          solveAsync(p, p->gradNew);
          launchVecOp(1, p->gradNew, nullptr, -1.0, n, nullptr, s);
          applyStep(p, p->gradNew, nullptr);
          vb_cost_stats statsAtSub{};
          const double newCostAtSub = evalCost(p, 1, &statsAtSub);
          const double secSub = since(startSub);
          const double costReductionAtSub = prevCost - newCostAtSub;
          const double ratioAtSub = costReductionAtSub / (modelCostReduction * appliedStepFactor);
          if (ratioAtSub >= st->min_relative_cost_reduction && acceptable(statsAtSub)) {
            newCost = newCostAtSub; stats = statsAtSub; costReduction = costReductionAtSub;
            ratioReductionToExpected = ratioAtSub; failureRateIsAcceptable = true;
            branch |= VB_BR_ACCEPT_AT_SUB;
            say(std::string("\\!/ cost reduction obtained applying factor ") + fbuf + " + sub-step:\n    solve: " + fmtDuration(secSub));
            break;
          }
        }
        if (!dontRetryFailed) { // Q5 latch
          dontRetryFailed = true;
          branch |= VB_BR_LATCH_NO_RETRY;
          say("\\!/ failing factors will no longer be retried!");
        }
        stepFactor = st->step_factor_decrease;
      }
    }
    const char* toleranceHit = ratioReductionToCost < st->relative_cost_tolerance ? "relative cost"
        : costReduction < st->absolute_cost_tolerance ? "absolute cost"
        : ratios[1] < st->variables_tolerance ? "variable" : nullptr;
    if (toleranceHit) branch |= VB_BR_TOLERANCE_HIT;
    if (st->verbose)
      std::printf("[viba_b200] it %d cost %.12g -> %.12g lambda %.3g model %.6g\n", iterationNum, prevCost, newCost, damping, modelCostReduction);
    const char* smiley;
    bool stopDamping = false;
    if (newCost > prevCost || !failureRateIsAcceptable) {
      if (lastTroubledIteration != iterationNum - 1) { troubledSeqStartDamping = damping; troubledSeqStart = iterationNum; }
      smiley = ":'(";
      branch |= VB_BR_REJECTED;
      damping *= st->damping_adjust_on_fail;
      VB_CUDA(cudaMemcpyAsync(p->dv.data, p->backup, p->varDataLen * sizeof(double), cudaMemcpyDeviceToDevice, s));
      if (damping > st->damping_max) { stopDamping = true; say("damping out of range, quadratic model failing?!"); }
      else lastTroubledIteration = iterationNum;
    } else {
      if (lastTroubledIteration == iterationNum - 1) {
        if (troubledSeqStartDamping < 1e1 && damping > 1e-3) {
          numTroubledSeqs++;
          largestTroubledSeq = std::max(largestTroubledSeq, iterationNum - troubledSeqStart);
          branch |= VB_BR_TROUBLED_SEQ_END;
        }
      }
      if (ratioReductionToExpected >= st->min_relative_cost_reduction && appliedStepFactor > st->min_step_factor_for_good) {
        smiley = toleranceHit ? ";-|" : ":-)";
        branch |= VB_BR_GOOD;
        damping *= st->damping_adjust_on_good;
        damping = std::max(damping, st->damping_min);
      } else {
        smiley = ":-/";
        branch |= VB_BR_AVERAGE;
        damping *= st->damping_adjust_on_average;
      }
      finalCost = newCost;
    }
    const double secTotal = since(startIt);
    if (wantTrace && hooks->trace_len < hooks->trace_cap) {
      vb_iter_trace& t = hooks->trace[hooks->trace_len++];
      t.prev_cost = prevCost; t.new_cost = newCost; t.damping = damping; t.damping_used = dampingUsed;
      t.model_cost_reduction = modelCostReduction; t.applied_step_factor = appliedStepFactor;
      t.ratio_reduction_to_expected = ratioReductionToExpected; t.cost_reduction = costReduction;
      t.grad_norm = gradNorm; t.step_norm = stepNorm;
      t.ratio_s2v_inf = ratios[0]; t.ratio_s2v_2 = ratios[1]; t.ratio_s2v_1 = ratios[2];
      t.ms_total = secTotal * 1e3; t.ms_grad_hess = secGradHess * 1e3; t.ms_factor = secFactor * 1e3; t.ms_solve = secSolve * 1e3;
      t.num_total = stats.num_total; t.num_invalid = stats.num_invalid; t.num_prev_invalid = stats.num_prev_invalid;
      t.branch = branch; t.step_factor_attempts = attempts;
    }
    if (stopDamping) { stopReason = 2; break; } // Optimizer.cpp:1027-1034 leaves before the iteration is counted
    iterationNum++;
    if (wantLog) { // Optimizer.cpp:1058-1076
      const double prevFailing = stats.num_prev_invalid / std::max((double)stats.num_total, 1.0);
      const double newFailing = stats.num_invalid / std::max((double)stats.num_total, 1.0);
      std::string t;
      t += std::string(" ") + smiley + " cost: " + fmtG(prevCost) + " -> " + fmtG(newCost) + " (" + fmtPercent(newCost / prevCost - 1.0, 2) +
           "), t: " + fmtDuration(secTotal) + "\n";
      t += "     n." + std::to_string(iterationNum) + "; g/H: " + fmtDuration(secGradHess) + ", solver:\n";
      if (p->solverType == VB_SOLVER_DIRECT) t += "    direct, factor: " + fmtDuration(secFactor) + ", solve: " + fmtDuration(secSolve) + "\n";
      else   // Optimizer.cpp:298-327 (the per-stage split of the reference's report is folded into the two totals)
        t += std::string("    semi-precond(") + solverName(p->solverType) + "), factor[elims] + init precond: " +
             fmtDuration(secFactor) + "\n    solve[elims] + reduced-PCG + backsubst: " + fmtDuration(secSolve) + " (iters: " +
             std::to_string(p->pcgLastIterations) + ", residual: " + fmtG(p->pcgLastResidual) + ")\n";
      t += "     lmbd: " + fmtG(damping) + ", relRed: " + fmtPercent(ratioReductionToExpected) + ", improv: " + fmtG(costReduction) +
           ", modelImprov: " + fmtG(modelCostReduction) + "\n";
      t += "    |G|: " + fmtG(gradNorm) + ", |S|: " + fmtG(stepNorm) + ", |s/v|_inf: " + fmtG(ratios[0]) + ", |_2: " + fmtG(ratios[1]) +
           ", |_1: " + fmtG(ratios[2]) + "\n";
      t += "    Failing factors: " + fmtG(prevFailing) + " (" + fmtPercent(prevFailing) + ") -> " + fmtG(newFailing) + " (" + fmtPercent(newFailing) + ")";
      say(t);
    }
    if (!toleranceHit) lastImprovementIteration = iterationNum;
    if (iterationNum >= lastImprovementIteration + st->stop_if_no_improvement_for &&
        iterationNum >= lastTroubledIteration + st->distance_from_troubled_iteration) {
      stopReason = 0;
      say(std::string(" >_< converged! (hit ") + (toleranceHit ? toleranceHit : "(null)") + " tolerance, for " +
          std::to_string(st->stop_if_no_improvement_for) + " iterations)");
      break;
    } else if (iterationNum >= st->max_num_iterations) {
      stopReason = 1;
      say(" X-| iteration limit reached! (" + std::to_string(st->max_num_iterations) + " iterations)");
      break;
    }
  }
  syncCheck(p);
  fillSummary();
  }
  catch (const std::domain_error& ex) { g_err = ex.what(); stopReason = 3; fillSummary(); return VB_ERR_NUMERIC; } // the variables stay at the current iterate
  catch (const std::invalid_argument& ex) { g_err = ex.what(); return VB_ERR_ARG; }
  catch (const std::exception& ex) { g_err = ex.what(); return VB_ERR_CUDA; }
  return VB_OK;
}
int vb_optimize(vb_problem* p, const vb_settings* st, vb_summary* out) { return vb_optimize_ex(p, st, nullptr, out); }

// one accepted-path LM iteration (Optimizer.cpp:800-1076 on the good-step path) with CUDA-event
// timing of the phases the reference logs (g/H, factor, solve: Optimizer.cpp:1062-1067).
int vb_lm_iteration_timed(vb_problem* p, double lambda, int apply, vb_iter_times* t, double* cost_before, double* cost_after) {
  VB_TRY
  requireProblem(p);
  cudaStream_t s = p->ctx->stream;
  cudaEvent_t* e = p->ev;
  const int64_t n = p->S.order;
  VB_CUDA(cudaEventRecord(e[0], s));
  launchPreint(p, -1);   // preStepCallback (Optimizer.cpp:801-805): part of the iteration, timed with grad/Hess
  launchRsTables(p, -1);
  evalGradHess(p, p->grad, 1, 1, 0, false); // also clears the evaluation error word
  VB_CUDA(cudaEventRecord(e[1], s));
  launchDamping(p->hess, p->diagOff, p->diagStride, p->diagSize, p->S.nB, lambda, p->dampMask, s);
  launchVecOp(0, p->step, p->grad, 0, n, nullptr, s);
  VB_CUDA(cudaEventRecord(e[2], s));
  factorAsync(p, e[3], p->step);
  VB_CUDA(cudaEventRecord(e[4], s));
  solveAsync(p, p->step);
  VB_CUDA(cudaMemsetAsync(p->acc + kAccDot, 0, sizeof(double), s));
  launchVecOp(2, p->step, p->grad, 0, n, p->acc + kAccDot, s);
  launchVecOp(1, p->step, nullptr, -1.0, n, nullptr, s);
  VB_CUDA(cudaEventRecord(e[5], s));
  VB_CUDA(cudaMemcpyAsync(p->backup, p->dv.data, p->varDataLen * sizeof(double), cudaMemcpyDeviceToDevice, s));
  applyStep(p, p->step, nullptr);
  VB_CUDA(cudaEventRecord(e[6], s));
  const size_t ns = p->ds.size();
  VB_CUDA(cudaMemsetAsync(p->acc + kAccStores, 0, 4 * ns * sizeof(double), s));
  for (size_t i = 0; i < ns; i++) launchCost(p->ds[i], p->dv, 1, p->acc + kAccStores + 4 * i, s);
  collective(p, VB_COLL_SUM, p->acc + kAccDot, 3 + 4 * (int64_t)ns); // step.grad, cost before, error word, cost stats
  VB_CUDA(cudaEventRecord(e[7], s));
  if (!apply) VB_CUDA(cudaMemcpyAsync(p->dv.data, p->backup, p->varDataLen * sizeof(double), cudaMemcpyDeviceToDevice, s));
  VB_CUDA(cudaMemcpyAsync(p->accHost, p->acc, (64 + 4 * ns) * sizeof(double), cudaMemcpyDeviceToHost, s));
  checkNumeric(p);
  checkRsTables(p, -1);
  checkPreint(p, -1);
  if (p->accHost[kAccErr] != 0.0) throw std::domain_error("RollingShutterData::getEstimate: time shift out of the sampled range");
  auto ms = [&](int a, int b) { float f = 0; cudaEventElapsedTime(&f, e[a], e[b]); return (double)f; };
  if (t) {
    t->grad_hess = ms(0, 1); t->damping = ms(1, 2); t->factor_schur = ms(2, 3); t->factor_chol = ms(3, 4);
    t->solve = ms(4, 5); t->apply = ms(5, 6); t->cost = ms(6, 7); t->total = ms(0, 7);
    t->step_dot_grad = p->accHost[kAccDot];
  }
  if (cost_before) *cost_before = p->accHost[kAccCost];
  if (cost_after) { double c = 0; for (size_t i = 0; i < ns; i++) c += p->accHost[kAccStores + 4 * i]; *cost_after = c; }
  VB_CATCH_NUMERIC
}

static int copyOut(vb_problem* p, const double* src, int64_t len, double* out, int64_t n) {
  VB_TRY
  requireProblem(p);
  if (!out || n != len) throw std::invalid_argument("buffer length mismatch");
  syncCheck(p);
  VB_CUDA(cudaMemcpy(out, src, len * sizeof(double), cudaMemcpyDeviceToHost));
  VB_CATCH
}
static int copyIn(vb_problem* p, double* dst, int64_t len, const double* in, int64_t n) {
  VB_TRY
  requireProblem(p);
  if (!in || n != len) throw std::invalid_argument("buffer length mismatch");
  syncCheck(p);
  VB_CUDA(cudaMemcpy(dst, in, len * sizeof(double), cudaMemcpyHostToDevice));
  VB_CATCH
}
int vb_download_vars(vb_problem* p, double* o, int64_t n) { return p ? copyOut(p, p->dv.data, p->varDataLen, o, n) : VB_ERR_ARG; }
int vb_upload_vars(vb_problem* p, const double* i, int64_t n) { return p ? copyIn(p, p->dv.data, p->varDataLen, i, n) : VB_ERR_ARG; }
int vb_download_grad(vb_problem* p, double* o, int64_t n) { return p ? copyOut(p, p->grad, p->S.order, o, n) : VB_ERR_ARG; }
int vb_download_step(vb_problem* p, double* o, int64_t n) { return p ? copyOut(p, p->step, p->S.order, o, n) : VB_ERR_ARG; }
int vb_download_hess(vb_problem* p, double* o, int64_t n) { return p ? copyOut(p, p->hess, p->S.dataSize, o, n) : VB_ERR_ARG; }
int vb_upload_hess(vb_problem* p, const double* i, int64_t n) { return p ? copyIn(p, p->hess, p->S.dataSize, i, n) : VB_ERR_ARG; }
int vb_upload_grad(vb_problem* p, const double* i, int64_t n) { return p ? copyIn(p, p->grad, p->S.order, i, n) : VB_ERR_ARG; }
int vb_download_factor_cache(vb_problem* p, int32_t store, double* o, int64_t n) {
  if (!p || store < 0 || store >= (int)p->ds.size()) { g_err = "store index out of range"; return VB_ERR_ARG; }
  return copyOut(p, p->ds[store].cache, p->ds[store].count, o, n);
}
int vb_upload_store_aux(vb_problem* p, int32_t store, const double* aux, int64_t n) {
  if (!p || store < 0 || store >= (int)p->ds.size()) { g_err = "store index out of range"; return VB_ERR_ARG; }
  return copyIn(p, const_cast<double*>(p->ds[store].aux), (int64_t)p->g.stores[store].aux.size(), aux, n);
}
int vb_upload_store_consts(vb_problem* p, int32_t store, const double* c, int64_t n) {
  if (!p || store < 0 || store >= (int)p->ds.size()) { g_err = "store index out of range"; return VB_ERR_ARG; }
  return copyIn(p, const_cast<double*>(p->ds[store].consts), (int64_t)p->g.stores[store].consts.size(), c, n);
}
int vb_download_store_aux(vb_problem* p, int32_t store, double* out, int64_t n) {
  if (!p || store < 0 || store >= (int)p->ds.size()) { g_err = "store index out of range"; return VB_ERR_ARG; }
  return copyOut(p, p->ds[store].aux, (int64_t)p->g.stores[store].aux.size(), out, n);
}
int vb_rs_set_imu(vb_problem* p, int32_t store, int64_t n_rigs, const int64_t* rig_mid_time_us, const int32_t* rig_calib_var,
                  int64_t half_length_us, int64_t n_samples, const int64_t* sample_time_ns, const float* gyro, const float* accel) {
  VB_TRY
  requireProblem(p);
  if (store < 0 || store >= (int)p->ds.size() || p->g.stores[store].type != VB_F_VISUAL_RS) throw std::invalid_argument("not a rolling-shutter store");
  if (n_samples == 0) { // forget the samples of this store: its tables stay as they are
    for (auto it = p->rsSources.begin(); it != p->rsSources.end();) it = it->store == store ? p->rsSources.erase(it) : it + 1;
    return VB_OK;
  }
  const auto& hs = p->g.stores[store];
  if (!rig_mid_time_us || !rig_calib_var || !sample_time_ns || !gyro || !accel || n_samples < 2 || half_length_us <= 0)
    throw std::invalid_argument("null or empty IMU data");
  if (hs.aux.size() < 4 || n_rigs != (int64_t)hs.aux[3] || (int64_t)hs.aux.size() != 4 + n_rigs * kRsTableStride)
    throw std::invalid_argument("rig count does not match the store's tables");
  for (int64_t i = 1; i < n_samples; i++)
    if (sample_time_ns[i] <= sample_time_ns[i - 1]) throw std::invalid_argument("IMU sample times must increase");
  std::vector<int64_t> calibOff((size_t)n_rigs, -1);
  for (int64_t r = 0; r < n_rigs; r++) {
    const int32_t v = rig_calib_var[r];
    if (v < 0) continue;
    if (v >= p->g.nv || p->g.kind[v] != VB_VAR_IMUCALIB) throw std::invalid_argument("rig_calib_var must name IMU calibration variables");
    calibOff[(size_t)r] = p->g.off[v];
  }
  VB_CUDA(cudaSetDevice(p->ctx->device));
  vb_problem::RsSource src;
  src.store = store; src.nRigs = n_rigs; src.nSamples = n_samples; src.halfUs = half_length_us;
  src.timeNs = p->upload(std::vector<int64_t>(sample_time_ns, sample_time_ns + n_samples));
  src.gyro = p->upload(std::vector<float>(gyro, gyro + 3 * n_samples));
  src.accel = p->upload(std::vector<float>(accel, accel + 3 * n_samples));
  src.midUs = p->upload(std::vector<int64_t>(rig_mid_time_us, rig_mid_time_us + n_rigs));
  src.calibOff = p->upload(calibOff);
  src.status = p->alloc<int>(1);
  for (auto it = p->rsSources.begin(); it != p->rsSources.end();)
    it = it->store == store ? p->rsSources.erase(it) : it + 1;
  p->rsSources.push_back(src);
  VB_CATCH
}
int vb_rs_set_gravity_var(vb_problem* p, int32_t store, int32_t gravity_var) {
  VB_TRY
  requireProblem(p);
  if (gravity_var >= 0 && (gravity_var >= p->g.nv || p->g.kind[gravity_var] != VB_VAR_S2)) throw std::invalid_argument("gravity_var must name an S2 variable");
  bool found = false;
  for (auto& src : p->rsSources)
    if (src.store == store) { src.gravOff = gravity_var < 0 ? -1 : p->g.off[gravity_var]; found = true; }
  if (!found) throw std::invalid_argument("store has no IMU samples (vb_rs_set_imu)");
  VB_CATCH
}
int vb_rs_refresh(vb_problem* p, int32_t store) {
  VB_TRY
  requireProblem(p);
  VB_CUDA(cudaSetDevice(p->ctx->device));
  refreshRsTables(p, store);
  VB_CATCH
}
// Optimizer::computeJointCovariances (Optimizer.cpp:503-612) for one list of registered variables: linearise at the
// current values (ResultCache untouched, :343-349), damp (:351-353), factor — with the reference's retry on a failed
// factorisation (:516-533: damping += 1e-9, then doubled) — and read the block of H^-1 column by column: one solve per
// tangent scalar of the listed variables (the reference restricts each solve to the trailing part of the factor;
// here the full forward / backward passes run, one right-hand side at a time).
// Optimizer::computeJointCovariances (Optimizer.cpp:503-612) for MANY lists at once, as SingleSessionProblem::computeCovariances calls
// it (one list per rig + one per calibration variable, SingleSessionProblem.cpp:100-126): linearise, damp and factor ONCE, then
// one solve per tangent scalar of every list.
int vb_compute_joint_covariances(vb_problem* p, double damping, int64_t n_lists, const int64_t* list_ptr, const int64_t* indices, double* cov,
                                 int64_t cov_len, int64_t* cov_ptr, double* damping_used) {
  VB_TRY
  requireProblem(p);
  if (p->nranks > 1) throw std::logic_error("vb_compute_joint_covariance: not available on a sharded problem");
  if (p->solverType != VB_SOLVER_DIRECT) throw std::logic_error("vb_compute_joint_covariance: needs the direct solver (vb_set_linear_solver)");
  if (n_lists <= 0 || !list_ptr || !indices || !cov) throw std::invalid_argument("null or empty index list");
  const Symbolic& S = p->S;
  // validate every list and lay the output out before any device work
  std::vector<int64_t> covOff((size_t)n_lists + 1, 0);
  for (int64_t l = 0; l < n_lists; l++) {
    const int64_t n = list_ptr[l + 1] - list_ptr[l];
    if (n <= 0) throw std::invalid_argument("null or empty index list");
    const int64_t* idx = indices + list_ptr[l];
    int64_t D = 0;
    for (int64_t i = 0; i < n; i++) {
      if (idx[i] < 0 || idx[i] >= S.nB) throw std::invalid_argument("taking index from an unregistered variable?");
      for (int64_t j = 0; j < i; j++) if (idx[j] == idx[i]) throw std::invalid_argument("duplicated indices");
      D += S.bsize[idx[i]];
    }
    covOff[(size_t)l + 1] = covOff[(size_t)l] + D * D;
  }
  if (cov_len != covOff[(size_t)n_lists]) throw std::invalid_argument("cov must hold (sum of tangent dims)^2 doubles per list");
  if (cov_ptr) for (int64_t l = 0; l <= n_lists; l++) cov_ptr[l] = covOff[(size_t)l];
  cudaStream_t s = p->ctx->stream;
  while (true) {
    evalGradHess(p, p->grad, 1, 0, 0, true);
    if (damping > 0) launchDamping(p->hess, p->diagOff, p->diagStride, p->diagSize, S.nB, damping, p->dampMask, s);
    factorAsync(p, nullptr);
    bool ok = true;
    try { checkNumeric(p); } catch (const std::domain_error&) { ok = false; }
    if (ok) break;
    if (damping < 1e-9) damping += 1e-9; else damping *= 2.0;
    if (damping > 1e12) throw std::domain_error("unable to compute covariances: the system stays singular");
  }
  if (damping_used) *damping_used = damping;
  const double one = 1.0;
  std::vector<int64_t> start, size, off;
  std::vector<double> col;
  for (int64_t l = 0; l < n_lists; l++) {
    const int64_t n = list_ptr[l + 1] - list_ptr[l];
    const int64_t* idx = indices + list_ptr[l];
    start.assign((size_t)n, 0); size.assign((size_t)n, 0); off.assign((size_t)n + 1, 0);
    for (int64_t i = 0; i < n; i++) {
      start[(size_t)i] = S.vecStart[S.pos[idx[i]]];
      size[(size_t)i] = S.bsize[idx[i]];
      off[(size_t)i + 1] = off[(size_t)i] + size[(size_t)i];
    }
    const int64_t D = off[(size_t)n];
    double* out = cov + covOff[(size_t)l];
    col.assign((size_t)D, 0.0);
    for (int64_t i = 0; i < n; i++)
      for (int64_t k = 0; k < size[(size_t)i]; k++) {
        VB_CUDA(cudaMemsetAsync(p->step, 0, S.order * sizeof(double), s));
        VB_CUDA(cudaMemcpyAsync(p->step + start[(size_t)i] + k, &one, sizeof(double), cudaMemcpyHostToDevice, s));
        p->partial[0] = false;
        solveAsync(p, p->step);
        for (int64_t j = 0; j < n; j++)
          VB_CUDA(cudaMemcpyAsync(col.data() + off[(size_t)j], p->step + start[(size_t)j], size[(size_t)j] * sizeof(double), cudaMemcpyDeviceToHost, s));
        syncCheck(p);
        for (int64_t r = 0; r < D; r++) out[r * D + off[(size_t)i] + k] = col[(size_t)r];
      }
  }
  VB_CATCH_NUMERIC
}
int vb_compute_joint_covariance(vb_problem* p, double damping, int64_t n, const int64_t* indices, double* cov, int64_t cov_len,
                                int64_t* offsets, double* damping_used) {
  if (offsets && p && indices && n > 0) {   // block offsets inside the one list (the batched call reports offsets between lists)
    int64_t o = 0;
    offsets[0] = 0;
    for (int64_t i = 0; i < n; i++) {
      if (indices[i] >= 0 && indices[i] < p->S.nB) o += p->S.bsize[indices[i]];
      offsets[i + 1] = o;
    }
  }
  const int64_t ptr[2] = {0, n};
  return vb_compute_joint_covariances(p, damping, 1, ptr, indices, cov, cov_len, nullptr, damping_used);
}
// Optimizer::computeMarginalProblem (Optimizer.cpp:422-494): H = Schur complement of every other variable onto the listed ones
// (with the damping of initDirectSolverData), b = the matching gradient, cost = cost - 0.5 |L^-1 g|^2 over the eliminated
// part.  The reference orders the listed variables last and reads the bottom-right block of a partial factorisation; the
// ordering here is fixed at problem creation, so the same quantities come from complete solves on the factored system:
// C = [(H + D)^-1]_ii column by column (as vb_compute_joint_covariance), x = (H + D)^-1 g, then H_m = C^-1, b_m = H_m x_i,
// g^T x = g_r^T H_rr^-1 g_r + b_m^T x_i.
int vb_compute_marginal_problem(vb_problem* p, double damping, int64_t n, const int64_t* indices, double* H, int64_t h_len, double* b,
                                int64_t b_len, double* cost, int64_t* offsets, double* damping_used) {
  VB_TRY
  requireProblem(p);
  if (p->nranks > 1) throw std::logic_error("vb_compute_marginal_problem: not available on a sharded problem");
  if (p->solverType != VB_SOLVER_DIRECT) throw std::logic_error("vb_compute_marginal_problem: needs the direct solver (vb_set_linear_solver)");
  if (n <= 0 || !indices || !H || !b) throw std::invalid_argument("null or empty index list");
  const Symbolic& S = p->S;
  std::vector<int64_t> start((size_t)n), size((size_t)n), off((size_t)n + 1, 0);
  for (int64_t i = 0; i < n; i++) {
    if (indices[i] < 0 || indices[i] >= S.nB) throw std::invalid_argument("taking index from an unregistered variable?");
    for (int64_t j = 0; j < i; j++) if (indices[j] == indices[i]) throw std::invalid_argument("duplicated indices");
    start[(size_t)i] = S.vecStart[S.pos[indices[i]]];
    size[(size_t)i] = S.bsize[indices[i]];
    off[(size_t)i + 1] = off[(size_t)i] + size[(size_t)i];
  }
  const int64_t D = off[(size_t)n];
  if (h_len != D * D || b_len != D) throw std::invalid_argument("H must hold D^2 and b D doubles, D = sum of the tangent dims");
  if (offsets) for (int64_t i = 0; i <= n; i++) offsets[i] = off[(size_t)i];
  cudaStream_t s = p->ctx->stream;
  double fullCost = 0.0;
  while (true) {
    fullCost = evalGradHess(p, p->grad, 1, 0, 0, true);
    if (damping > 0) launchDamping(p->hess, p->diagOff, p->diagStride, p->diagSize, S.nB, damping, p->dampMask, s);
    factorAsync(p, nullptr);
    bool ok = true;
    try { checkNumeric(p); } catch (const std::domain_error&) { ok = false; }
    if (ok) break;
    if (damping < 1e-9) damping += 1e-9; else damping *= 2.0;
    if (damping > 1e12) throw std::domain_error("unable to compute the marginal problem: the system stays singular");
  }
  if (damping_used) *damping_used = damping;
  std::vector<double> C((size_t)(D * D)), col((size_t)D), x((size_t)D);
  const double one = 1.0;
  for (int64_t i = 0; i < n; i++)
    for (int64_t k = 0; k < size[(size_t)i]; k++) {
      VB_CUDA(cudaMemsetAsync(p->step, 0, S.order * sizeof(double), s));
      VB_CUDA(cudaMemcpyAsync(p->step + start[(size_t)i] + k, &one, sizeof(double), cudaMemcpyHostToDevice, s));
      p->partial[0] = false;
      solveAsync(p, p->step);
      for (int64_t j = 0; j < n; j++)
        VB_CUDA(cudaMemcpyAsync(col.data() + off[(size_t)j], p->step + start[(size_t)j], size[(size_t)j] * sizeof(double), cudaMemcpyDeviceToHost, s));
      syncCheck(p);
      for (int64_t r = 0; r < D; r++) C[(size_t)(r * D + off[(size_t)i] + k)] = col[(size_t)r];
    }
  // x = (H + D)^-1 g and g^T x
  launchVecOp(0, p->step, p->grad, 0, S.order, nullptr, s);
  p->partial[0] = false;
  solveAsync(p, p->step);
  const double gTx = dotVec(p, p->step, p->grad);
  for (int64_t j = 0; j < n; j++)
    VB_CUDA(cudaMemcpyAsync(x.data() + off[(size_t)j], p->step + start[(size_t)j], size[(size_t)j] * sizeof(double), cudaMemcpyDeviceToHost, s));
  syncCheck(p);
  // H_m = C^-1 on the host: C is symmetric positive definite (symmetrised first: its columns come from separate solves)
  std::vector<double> L((size_t)(D * D), 0.0);
  for (int64_t r = 0; r < D; r++)
    for (int64_t c = 0; c <= r; c++) L[(size_t)(r * D + c)] = 0.5 * (C[(size_t)(r * D + c)] + C[(size_t)(c * D + r)]);
  for (int64_t j = 0; j < D; j++) {
    double d = L[(size_t)(j * D + j)];
    for (int64_t k = 0; k < j; k++) d -= L[(size_t)(j * D + k)] * L[(size_t)(j * D + k)];
    if (!(d > 0.0)) throw std::domain_error("vb_compute_marginal_problem: the covariance block is not positive definite");
    const double ljj = std::sqrt(d);
    L[(size_t)(j * D + j)] = ljj;
    for (int64_t i = j + 1; i < D; i++) {
      double t = L[(size_t)(i * D + j)];
      for (int64_t k = 0; k < j; k++) t -= L[(size_t)(i * D + k)] * L[(size_t)(j * D + k)];
      L[(size_t)(i * D + j)] = t / ljj;
    }
  }
  std::vector<double> e((size_t)D);
  for (int64_t c = 0; c < D; c++) {   // column c of C^-1: L L^T h = e_c
    for (int64_t i = 0; i < D; i++) {
      double t = i == c ? 1.0 : 0.0;
      for (int64_t k = 0; k < i; k++) t -= L[(size_t)(i * D + k)] * e[(size_t)k];
      e[(size_t)i] = t / L[(size_t)(i * D + i)];
    }
    for (int64_t i = D - 1; i >= 0; i--) {
      double t = e[(size_t)i];
      for (int64_t k = i + 1; k < D; k++) t -= L[(size_t)(k * D + i)] * e[(size_t)k];
      e[(size_t)i] = t / L[(size_t)(i * D + i)];
    }
    for (int64_t r = 0; r < D; r++) H[r * D + c] = e[(size_t)r];
  }
  double bTx = 0.0;
  for (int64_t r = 0; r < D; r++) {
    double t = 0.0;
    for (int64_t c = 0; c < D; c++) t += H[r * D + c] * x[(size_t)c];
    b[r] = t;
    bTx += t * x[(size_t)r];
  }
  if (cost) *cost = fullCost - 0.5 * (gTx - bTx);
  VB_CATCH_NUMERIC
}
int vb_refine_points(vb_problem* p, vb_refine_stats* out, double* per_point, int64_t n_points) {
  VB_TRY
  requireProblem(p);
  if (p->nranks > 1) throw std::logic_error("vb_refine_points: not available on a sharded problem");
  if (!p->refineReady) throw std::logic_error("vb_refine_points: the problem has no visual factors in the packed layout");
  if (per_point && n_points != p->refine.nLm) throw std::invalid_argument("per_point must hold 3 doubles per refined landmark");
  VB_CUDA(cudaSetDevice(p->ctx->device));
  cudaStream_t s = p->ctx->stream;
  VB_CUDA(cudaMemsetAsync(p->acc, 0, 8 * sizeof(double), s));
  VB_CUDA(cudaMemsetAsync(p->acc + kAccErrSlot, 0, sizeof(double), s));
  launchRefinePoints(p->refine, p->dv, p->refinePerPoint, p->acc, s);
  VB_CUDA(cudaMemcpyAsync(p->accHost, p->acc, 8 * sizeof(double), cudaMemcpyDeviceToHost, s));
  VB_CUDA(cudaMemcpyAsync(p->accHost + kAccErrSlot, p->acc + kAccErrSlot, sizeof(double), cudaMemcpyDeviceToHost, s));
  if (per_point) VB_CUDA(cudaMemcpyAsync(per_point, p->refinePerPoint, 3 * p->refine.nLm * sizeof(double), cudaMemcpyDeviceToHost, s));
  syncCheck(p);
  if (p->accHost[kAccErrSlot] != 0.0) throw std::domain_error("RollingShutterData::getEstimate: time shift outside the sampled window");
  if (out) {
    out->total_start_cost = p->accHost[0]; out->total_end_cost = p->accHost[1];
    out->num_points = (int64_t)p->accHost[2]; out->num_failures = (int64_t)p->accHost[3];
    out->num_total_iterations = (int64_t)p->accHost[4]; out->num_at_least_one_iteration = (int64_t)p->accHost[5];
  }
  VB_CATCH
}
int64_t vb_refine_num_points(const vb_problem* p) { return p && p->refineReady ? p->refine.nLm : 0; }
int vb_refine_triangulation(vb_problem* p, double outlier_threshold, int32_t skip_outliers, int32_t max_num_iterations, double huber_radius,
                            double* per_point, int64_t n_points, int64_t* num_failures) {
  VB_TRY
  requireProblem(p);
  if (p->nranks > 1) throw std::logic_error("vb_refine_triangulation: not available on a sharded problem");
  if (!p->refineReady) throw std::logic_error("vb_refine_triangulation: the problem has no visual factors in the packed layout");
  if (per_point && n_points != p->refine.nLm) throw std::invalid_argument("per_point must hold 3 doubles per refined landmark");
  if (!(outlier_threshold > 0.0) || !(huber_radius > 0.0) || max_num_iterations < 0)   // HuberLoss: BASPACHO_CHECK_GT(a, 0) (SoftLoss.h:66-68)
    throw std::invalid_argument("vb_refine_triangulation: outlier_threshold > 0, huber_radius > 0, max_num_iterations >= 0 required");
  cudaStream_t s = p->ctx->stream;
  VB_CUDA(cudaMemsetAsync(p->acc, 0, 8 * sizeof(double), s));
  VB_CUDA(cudaMemsetAsync(p->acc + kAccErrSlot, 0, sizeof(double), s));
  launchRefineTriangulation(p->refine, p->dv, outlier_threshold, skip_outliers != 0, max_num_iterations, huber_radius, p->refinePerPoint, p->acc, s);
  VB_CUDA(cudaMemcpyAsync(p->accHost, p->acc, 8 * sizeof(double), cudaMemcpyDeviceToHost, s));
  VB_CUDA(cudaMemcpyAsync(p->accHost + kAccErrSlot, p->acc + kAccErrSlot, sizeof(double), cudaMemcpyDeviceToHost, s));
  if (per_point) VB_CUDA(cudaMemcpyAsync(per_point, p->refinePerPoint, 3 * p->refine.nLm * sizeof(double), cudaMemcpyDeviceToHost, s));
  syncCheck(p);
  if (p->accHost[kAccErrSlot] != 0.0) throw std::domain_error("RollingShutterData::getEstimate: time shift outside the sampled window");
  if (num_failures) *num_failures = (int64_t)p->accHost[0];
  VB_CATCH_NUMERIC
}
int vb_refine_triangulation_inliers(vb_problem* p, int32_t store, uint8_t* flags, int64_t n_factors) {
  VB_TRY
  requireProblem(p);
  if (!p->refineReady) throw std::logic_error("vb_refine_triangulation_inliers: the problem has no visual factors in the packed layout");
  int rsIdx = -1;
  for (size_t k = 0; k < p->refineStoreId.size(); k++) if (p->refineStoreId[k] == store) rsIdx = (int)k;
  if (rsIdx < 0) throw std::invalid_argument("vb_refine_triangulation_inliers: not a visual store");
  if (!flags || n_factors != (int64_t)p->g.stores[(size_t)store].count) throw std::invalid_argument("flags must hold one byte per factor of the store");
  const std::vector<int32_t>& slot = p->refineSlotFactor[(size_t)rsIdx];
  std::vector<double> host(slot.size());
  cudaStream_t s = p->ctx->stream;
  if (!slot.empty()) VB_CUDA(cudaMemcpyAsync(host.data(), p->refine.st[rsIdx].backup, slot.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
  syncCheck(p);
  std::memset(flags, 0, (size_t)n_factors);
  for (size_t i = 0; i < slot.size(); i++) flags[slot[i]] = host[i] == 1.0 ? 1 : 0;
  VB_CATCH
}
int vb_preint_set_imu(vb_problem* p, int32_t store, int64_t n_factors, const int64_t* start_us, const int64_t* end_us, const int32_t* calib_var,
                      int64_t n_samples, const int64_t* sample_time_ns, const float* gyro, const float* accel, const double* gyro_var3,
                      const double* accel_var3) {
  VB_TRY
  requireProblem(p);
  if (store < 0 || store >= (int)p->ds.size()) throw std::invalid_argument("store index out of range");
  const int type = p->g.stores[store].type;
  if (type != VB_F_INERTIAL && type != VB_F_INERTIAL_SEC && type != VB_F_INERTIAL_SEC_SPLIT) throw std::invalid_argument("not an inertial store");
  auto forget = [&]() { for (auto it = p->preintSources.begin(); it != p->preintSources.end();) it = it->store == store ? p->preintSources.erase(it) : it + 1; };
  if (n_samples == 0) { forget(); return VB_OK; }
  const auto& hs = p->g.stores[store];
  if (!start_us || !end_us || !calib_var || !sample_time_ns || !gyro || !accel || !gyro_var3 || !accel_var3 || n_samples < 2)
    throw std::invalid_argument("null or empty IMU data");
  if (n_factors != hs.count || (int64_t)hs.aux.size() != n_factors * kPreintStride) throw std::invalid_argument("factor count does not match the store's records");
  for (int64_t i = 1; i < n_samples; i++)
    if (sample_time_ns[i] <= sample_time_ns[i - 1]) throw std::invalid_argument("IMU sample times must increase");
  std::vector<int64_t> calibOff((size_t)n_factors, -1);
  std::vector<uint32_t> calibFlags((size_t)n_factors, 0);
  for (int64_t k = 0; k < n_factors; k++) {
    const int32_t v = calib_var[k];
    if (v < 0) continue;
    if (v >= p->g.nv || p->g.kind[v] != VB_VAR_IMUCALIB) throw std::invalid_argument("calib_var must name IMU calibration variables");
    if (!(start_us[k] < end_us[k])) throw std::invalid_argument("empty preintegration interval");
    calibOff[(size_t)k] = p->g.off[v];
    calibFlags[(size_t)k] = p->g.flags[v];
  }
  VB_CUDA(cudaSetDevice(p->ctx->device));
  vb_problem::PreintSource src;
  src.store = store; src.nFactors = n_factors; src.nSamples = n_samples;
  src.timeNs = p->upload(std::vector<int64_t>(sample_time_ns, sample_time_ns + n_samples));
  src.gyro = p->upload(std::vector<float>(gyro, gyro + 3 * n_samples));
  src.accel = p->upload(std::vector<float>(accel, accel + 3 * n_samples));
  src.startUs = p->upload(std::vector<int64_t>(start_us, start_us + n_factors));
  src.endUs = p->upload(std::vector<int64_t>(end_us, end_us + n_factors));
  src.calibOff = p->upload(calibOff);
  src.calibFlags = p->upload(calibFlags);
  src.noise6 = p->upload(std::vector<double>{gyro_var3[0], gyro_var3[1], gyro_var3[2], accel_var3[0], accel_var3[1], accel_var3[2]});
  src.cov = p->alloc<double>((size_t)n_factors * 81);
  src.omegaAtEnd = p->alloc<double>((size_t)n_factors * 3);
  VB_CUDA(cudaMemset(src.cov, 0, (size_t)n_factors * 81 * sizeof(double)));
  VB_CUDA(cudaMemset(src.omegaAtEnd, 0, (size_t)n_factors * 3 * sizeof(double)));
  src.status = p->alloc<int>(1);
  forget();
  p->preintSources.push_back(src);
  VB_CATCH
}
int vb_preint_refresh(vb_problem* p, int32_t store) {
  VB_TRY
  requireProblem(p);
  VB_CUDA(cudaSetDevice(p->ctx->device));
  if (!launchPreint(p, store) && store >= 0) throw std::invalid_argument("store has no IMU samples (vb_preint_set_imu)");
  checkPreint(p, store);
  VB_CATCH
}
int vb_preint_download(vb_problem* p, int32_t store, double* cov, double* omega_at_end, int64_t n_factors) {
  VB_TRY
  requireProblem(p);
  for (auto& src : p->preintSources) {
    if (src.store != store) continue;
    if (n_factors != src.nFactors) throw std::invalid_argument("factor count mismatch");
    VB_CUDA(cudaStreamSynchronize(p->ctx->stream));
    if (cov) VB_CUDA(cudaMemcpy(cov, src.cov, (size_t)n_factors * 81 * sizeof(double), cudaMemcpyDeviceToHost));
    if (omega_at_end) VB_CUDA(cudaMemcpy(omega_at_end, src.omegaAtEnd, (size_t)n_factors * 3 * sizeof(double), cudaMemcpyDeviceToHost));
    return VB_OK;
  }
  throw std::invalid_argument("store has no IMU samples (vb_preint_set_imu)");
  VB_CATCH
}
int vb_measure_dmma_peak(vb_ctx* ctx, double* tflops) {
  VB_TRY
  if (!ctx || !tflops) throw std::invalid_argument("null argument");
  VB_CUDA(cudaSetDevice(ctx->device));
  *tflops = measureDmmaPeak(ctx->stream);
  VB_CATCH
}
int vb_measure_dmma_shapes(vb_ctx* ctx, double* tflops4) {
  VB_TRY
  if (!ctx || !tflops4) throw std::invalid_argument("null argument");
  VB_CUDA(cudaSetDevice(ctx->device));
  measureDmmaPeak(ctx->stream, tflops4);
  VB_CATCH
}

} // extern "C"
