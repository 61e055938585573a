// Internal definitions shared by the translation units of libviba_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <string>
#include <vector>
#include <stdexcept>
#include "symbolic.hpp"

namespace vb {

#define VB_CUDA(x)                                                                          \
  do {                                                                                      \
    cudaError_t e_ = (x);                                                                   \
    if (e_ != cudaSuccess) throw std::runtime_error(std::string(#x) + ": " + cudaGetErrorString(e_)); \
  } while (0)

constexpr int kPreintStride = 250; // R(4) dV(3) dP(3) dt(1) J(9x23) evalPoint(32)
constexpr int kMaxVars = 10;

struct LevelTask { int32_t lump, ti, tj; };
struct SchurTask { int32_t group, ti, tj; uint32_t mask; }; // mask: member landmarks present in both row tiles

// visual stores: observation tables packed in the two processing orders of kernels_visual.cu
struct VisPack {
  const int4* chunkHdr;    // [2 nChunks] {pose, extr, intr, vel data offsets}, {camFlags, camDim, first factor, 0}
  const int32_t* cLm;      // [nActive] chunk order: data offset of the landmark
  const double* cConst;    // [nActive * cpf] chunk order
  const int32_t* lVar;     // [nActive * 6] landmark-major order: data offsets {X, pose, extr, intr, vel}, camFlags
  const double* lConst;    // [nActive * cpf] landmark-major order
  const uint32_t* lOut;    // [nActive * 6] {g_p, H_pp, W_pose, W_extr, W_intr, W_vel} offsets, 0xffffffff = absent
};

// refinePoints (kernels_visual.cu): the visual stores seen landmark by landmark
struct RefineStore {
  const int32_t* lVar; const double* lConst; const int64_t* lmPtr;   // [nLm + 1] range of landmark j in the store's landmark-major tables
  double* backup;                                                    // [nActive] costBackups of PointRefinement.cpp:49-86
  const double* aux; double loss_a, loss_k; int32_t cpf, rs;
};
struct RefineArgs {
  RefineStore st[4]; int32_t nStores; int64_t nLm;
  const int64_t* lmOff;                                              // [nLm] data offset of landmark j
};

// device view of one factor store (SoA tables + scatter map)
struct DevStore {
  int32_t type, nvars, cpf, npairs;
  int64_t count;
  const int32_t* vars;   // [count*nvars] variable ids
  const double* consts;  // [count*cpf]
  const double* aux;     // store-level (preintegration records)
  double* cache;         // ResultCache (Factor.h:92-94): last valid 0.5*rho or -1
  double loss_a, loss_k;
  const int64_t* gOff;    // [count*nvars] scalar offset in the permuted vectors, -1 for constants
  const int64_t* hOff;    // [count*npairs] pair (i,j<=i) at i*(i+1)/2+j ; -1 if absent
  const int32_t* hStride; // [count*npairs] row stride | (transposed << 30)
  // visual stores: processing order grouped by (rig pose, extrinsics, intrinsics) and cut into
  // chunks of <= 128 observations that share all three camera-side variables
  const int32_t* order;      // [count] factor index
  const int32_t* chunkStart; // [nChunks+1] into order
  int32_t nChunks;
  int32_t twoPass;           // 1: camera side per chunk + landmark side in landmark-major order (two kernels)
  const int32_t* orderLm;    // [nActive] visual stores: owned factors sorted by (landmark, extrinsics, intrinsics)
  // shard view (vb_problem_create_sharded): the factors this rank evaluates.  Small stores: the
  // contiguous range [kBegin, kBegin + nActive); visual stores: the sorted list `active` (factors
  // whose landmark this rank owns).  Unsharded: kBegin = 0, nActive = count, active = null.
  int64_t kBegin, nActive;
  const int32_t* active;
  VisPack pk;
  int32_t visImpl;           // 1: kernels_visual.cu (packed tables), 0: the round-1 kernels of kernels_factors.cu
};

// device view of the variables
struct DevVars {
  int64_t nv;
  const int32_t* kind;
  const int32_t* tdim;
  const uint32_t* flags;
  const int64_t* off;    // offset of the variable's data
  const int64_t* vecOff; // permuted scalar offset or -1
  const double* aux;
  double* data;
  double* errAcc;        // set to 1.0 by a factor kernel on a hard evaluation error (RS time shift out of range)
};

// device view of the solver structure
struct DevSolver {
  int64_t nP, nC, nL;
  const int64_t* lmOff; const int64_t* lmRowPtr; const int32_t* lmRowCam; const int32_t* lmRowOff;
  const int32_t* camLump; const int32_t* camStart; const int32_t* lumpFirst;
  const int32_t* lumpWidth; const int32_t* lumpRows; const int64_t* lumpOff;
  const int64_t* lumpRowPtr; const int32_t* lumpRowCam; const int32_t* lumpRowOff;
  const int64_t* diagLOff;  // [nL] offset of the factored diagonal block in diagL
  double* diagL;
  double* diagInv;          // [reduced order x 32] inverses of the 32x32 diagonal blocks of every L
  // update tables (per lump, indexed by below-row scalar q)
  const int64_t* updPtr;    // [nL+1] start of the lump's below rows in the per-row arrays
  const int64_t* updColOff; // target address of column role: lumpOff[T] + col
  const int32_t* updStride; // width of T
  const int32_t* updTgt;    // target index t among the lump's distinct targets
  const int32_t* updGlobal; // reduced-system scalar index of the row
  const int64_t* relPtr;    // [nL+1] start of rel table (nT * r ints)
  const int32_t* rel;       // rel[t*r + q] = row of q inside T's panel
  int64_t camVec0;          // scalar offset of the camera section in the permuted vectors (= 3 nP)
  double* solveAcc;         // [reduced order] scratch of the backward solve (kept zero between calls)
  int* solveCnt;            // [nL] arrival counters of the solve kernels (kept zero between calls)
  // landmark groups of the Schur update (scatter source index nL + g in the upd*/rel tables)
  int64_t nG;
  const int32_t* grpChunks;  // [nG] chunks of <= 32 member landmarks
  const int64_t* grpBlkPtr;  // [nG+1] first row of the group's offset table (rows of 32 ints)
  const int32_t* grpBlkOff;  // [(grpBlkPtr[g] + c * nblk + b) * 32 + l] offset of member l of chunk c at union block b, or -1
  const int64_t* grpMaskPtr; // [nG+1] into grpTileMask
  const uint32_t* grpTileMask; // [grpMaskPtr[g] + c * ntiles + t] members of chunk c with rows in 64-row tile t
  const int32_t* grpRowBlk;  // per union scalar row (index updPtr[nL+g] - grpRowBase + q): union block
  const int32_t* grpRowSub;  //   ... and the row inside that block
  const int32_t* grpRows;    // [nG] union scalar rows
  const int32_t* grpBlks;    // [nG] union row blocks
  int64_t grpRowBase;        // updPtr[nL]
  // shard view: landmarks this rank owns (null = all nP)
  const int32_t* lmActive; int64_t nLmActive;
};


// kernels_factors.cu
void launchGradHess(const DevStore& st, const DevVars& v, double* grad, double* hess, int updateCache, int dontRetry,
                    double* costAcc, cudaStream_t s);
void launchCost(const DevStore& st, const DevVars& v, int comparable, double* acc /* cost,total,invalid,prevInvalid */, cudaStream_t s);
void launchApplyStep(const DevVars& v, const int64_t* blockVar, int64_t nB, const double* step, double* ratioAcc, cudaStream_t s);
void launchDamping(double* hess, const int64_t* diagOff, const int32_t* diagStride, const int32_t* diagSize, int64_t nB,
                   double lambda, const double* addMask, cudaStream_t s);
// kernels_visual.cu
void launchVisualGradHess(const DevStore& st, const DevVars& v, double* grad, double* hess, int updateCache, int dontRetry,
                          double* costAcc, cudaStream_t s);
void launchVisualCost(const DevStore& st, const DevVars& v, int comparable, double* acc, cudaStream_t s);
void launchRefinePoints(const RefineArgs& A, const DevVars& v, double* perPoint, double* statsAcc, cudaStream_t s);
void launchRefineTriangulation(const RefineArgs& A, const DevVars& v, double outlierThreshold, bool skipOutliers, int maxIts, double radius,
                               double* perPoint, double* statsAcc, cudaStream_t s);
// kernels_solver.cu
void launchSchurPrep(const DevSolver& d, double* hess, int* errFlag, cudaStream_t s);
void launchZeroOwnedPanels(const DevSolver& d, double* hess, cudaStream_t s);
void launchSchur(const DevSolver& d, double* hess, const SchurTask* tasks, int nTasks, int* errFlag, cudaStream_t s);
// rhs (nullable): the forward solve of that vector is fused into the factorisation (its lump segments ride through the trsm,
// the panel products through the update kernel's gemv tasks)
void launchLumpFactor(const DevSolver& d, double* hess, const int32_t* lumps, int nLumps, int maxRows, int maxWidth, int* errFlag, cudaStream_t s, double* rhs = nullptr);
void launchLumpUpdate(const DevSolver& d, double* hess, const LevelTask* tasks, int nTasks, int maxWidth, int useAtomics, cudaStream_t s, double* rhs = nullptr);
void launchSchurUpdate(const DevSolver& d, double* hess, const SchurTask* tasks, int nTasks, cudaStream_t s);
void launchSolveLandmarksFwd(const DevSolver& d, const double* hess, double* vec, cudaStream_t s);
void launchSolveLandmarksBwd(const DevSolver& d, const double* hess, double* vec, cudaStream_t s);
void launchSolveLumpsFwd(const DevSolver& d, const double* hess, double* vec, const int32_t* lumps, int nLumps, int maxWidth, int maxBelow, int useAtomics, cudaStream_t s);
void launchSolveLumpsBwd(const DevSolver& d, const double* hess, double* vec, const int32_t* lumps, int nLumps, int maxWidth, int maxBelow, cudaStream_t s);
// op: 0 a=b ; 1 a*=s ; 2 out[0]+=a.b ; 3 a=0
void launchVecOp(int op, double* a, const double* b, double s, int64_t n, double* out, cudaStream_t st);
// RollingShutterData::compute per rig (kernels_refresh.cu); status receives the largest rf::k* code
void launchRsRefresh(const int64_t* timeNs, const float* gyro, const float* accel, int64_t nSamples, const double* varData,
                     const int64_t* calibOff, const int64_t* midUs, int64_t halfUs, int64_t nRigs, double* tables, int* status,
                     cudaStream_t st);
void launchPreintRefresh(const int64_t* timeNs, const float* gyro, const float* accel, int64_t nSamples, const double* varData,
                         const int64_t* calibOff, const uint32_t* calibFlags, const int64_t* startUs, const int64_t* endUs, int64_t nFactors,
                         const double* noise6, double* scratch, double* records, double* cov, double* omegaAtEnd, int* status, cudaStream_t st);
// kernels_pcg.cu (iterative path)
int camSymMvRows();
void launchCamSymMv(const DevSolver& d, const double* hess, const int2* tasks, int nTasks, const double* v, double* u, int maxWidth, cudaStream_t s);
void launchJacobiInit(const DevSolver& d, const double* hess, const int64_t* jacOff, double* jac, int* errFlag, cudaStream_t s);
void launchJacobiApply(const DevSolver& d, const int64_t* jacOff, const double* jac, const double* in, double* out, cudaStream_t s);
void launchRoundToFloat(double* a, int64_t n, cudaStream_t s);
void launchCamDiagRaise(const DevSolver& d, double* hess, double eps, cudaStream_t s);
void launchPcgDot2(const double* a, const double* b, const double* c, const double* dd, int64_t n, double* out, cudaStream_t s);
void launchPcgStep(double* x, double* r, const double* p, const double* Ap, const double* num, const double* den, int64_t n, cudaStream_t s);
void launchPcgDir(double* p, const double* z, const double* num, const double* den, int64_t n, cudaStream_t s);
double measureDmmaPeak(cudaStream_t s, double* perShape = nullptr);

} // namespace vb
