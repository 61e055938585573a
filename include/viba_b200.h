/*
 * viba_b200.h — C ABI of the B200-native VIBA inner Levenberg–Marquardt iteration.
 *
 * This is the drop-in boundary.  Every entry point replaces a piece of the
 * reference's lib/small_thing runtime or of BaSpaCho as consumed by it; the reference file:line
 * each one replaces is cited beside it.  Plain C: pointers + sizes, no C++/torch types.
 * All functions return 0 on success, a VB_ERR_* code otherwise; vb_last_error() gives the text.
 * There is no CPU fallback: every compute entry point fails with VB_ERR_CUDA when no sm_100
 * device is usable.
 */
#ifndef VIBA_B200_H_
#define VIBA_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VB_OK 0
#define VB_ERR_ARG 1
#define VB_ERR_CUDA 2
#define VB_ERR_NUMERIC 3  /* non-SPD pivot, RS time shift out of table range, ... */
#define VB_ERR_STATE 4

/* Variable kinds (reference lib/small_thing/Variable.h:25-221,
 * interfaces/ark/camera_model/CameraModelParam.h:124-149, interfaces/ark/imu_model/ImuCalibParam.h:23) */
enum {
  VB_VAR_VEC3 = 0,     /* landmark / velocity / omega: 3 data, 3 tangent, += */
  VB_VAR_SE3 = 1,      /* 7 data = quat(x,y,z,w) + t ; 6 tangent [trans, rot] ; exp(d)*T */
  VB_VAR_S2 = 2,       /* 3 data (+radius in var_aux) ; 2 tangent */
  VB_VAR_CAM = 3,      /* 17 data = 15 Fisheye624 params, readout s, time offset s ; tangent 15..17 */
  VB_VAR_IMUCALIB = 4, /* 32 data ; tangent <= 23 */
  VB_VAR_VEC2 = 5
};

/* var_flags bits */
#define VB_CAM_EST_READOUT 1u
#define VB_CAM_EST_TIMEOFFSET 2u
#define VB_CAM_HAS_READOUT 4u

/* Factor types (one per FactorStore instantiation reachable from ark_vi_ba) */
enum {
  VB_F_VISUAL = 1,           /* B1  viba/problem/VisualFactor.cpp:36-82      vars: X,T_bw,T_cb,cam */
  VB_F_VISUAL_RS = 2,        /* B2  VisualFactor.cpp:122-214                 vars: + vel_world */
  VB_F_INERTIAL = 4,         /* B4  InertialFactor.cpp:19-127                calib,Tp,vp,Tn,vn,g */
  VB_F_INERTIAL_SEC = 5,     /* B5  InertialFactor.cpp:256-301 (common extrinsics) */
  VB_F_INERTIAL_SEC_SPLIT = 6, /* B6 InertialFactor.cpp:190-254 */
  VB_F_RW_IMUCALIB = 7,      /* B7  RandomWalkFactor.cpp:16-57 */
  VB_F_RW_CAM = 8,           /* B8  RandomWalkFactor.cpp:59-100 */
  VB_F_RW_SE3 = 9,           /* B9/B10 RandomWalkFactor.cpp:102-168 */
  VB_F_PRIOR_IMUCALIB = 11,  /* B11 PriorFactor.cpp:82-109 */
  VB_F_PRIOR_CAM = 12,       /* B12 PriorFactor.cpp:111-138 */
  VB_F_PRIOR_SE3 = 13,       /* B13/B14 PriorFactor.cpp:140-176 */
  VB_F_OMEGA_PRIOR = 15,     /* B15 OmegaPriorFactor.cpp:24-31 */
  VB_F_OMEGA_PRIOR_SEC = 16  /* B16 OmegaPriorFactor.cpp:35-53 */
};

/* One factor store = one FactorStore<...> instantiation (lib/small_thing/Factor.h:96-141).
 * Factors are given in insertion order; `vars` are ids into the variable pool in functor
 * argument order; `consts` is the per-factor payload (layout per type in DESIGN.md §3). */
typedef struct vb_factor_store {
  int32_t type;              /* VB_F_* */
  int32_t nvars;             /* variables per factor */
  int64_t count;
  const int32_t* vars;       /* [count * nvars] */
  int32_t consts_per_factor; /* doubles */
  const double* consts;      /* [count * consts_per_factor] */
  double loss_a, loss_k;     /* HuberLossWithCutoff(a,k) (SoftLoss.h:115-163); a<=0 => trivial loss */
  int64_t aux_len;           /* doubles of store-level aux data (preintegration / RS tables) */
  const double* aux;
} vb_factor_store;

/* The whole graph.  Registration order (var_index), factor order and store order are explicit
 * inputs because the reference derives them from libstdc++ unordered_map iteration
 * (lib/small_thing/Optimizer.cpp:25-39, Factor.h:737-785). */
typedef struct vb_graph_desc {
  int64_t num_vars;
  const int32_t* var_kind;   /* VB_VAR_* */
  const int32_t* var_tdim;   /* tangent dim */
  const uint32_t* var_flags;
  const int64_t* var_index;  /* param-block (registration) index, or -2 for constant (Variable.h:224-250) */
  const int64_t* var_off;    /* offset of the variable's data in var_data (doubles) */
  const double* var_aux;     /* one double per variable (S2 radius), may be NULL */
  int64_t var_data_len;
  const double* var_data;
  int64_t num_points;        /* elimRanges = [0, num_points) in registration-index space
                                (interfaces/ark/main_AriaKit_ViBa.cpp:76-77) */
  int32_t num_stores;
  const vb_factor_store* stores;
} vb_graph_desc;

/* LM settings: lib/small_thing/Optimizer.h:40-78 with the ark_vi_ba overrides
 * (viba/common/Settings.cpp:313-320).  vb_settings_default() fills the library defaults. */
typedef struct vb_settings {
  double damping, damping_adjust_on_fail, damping_adjust_on_good, damping_adjust_on_average;
  double damping_min, damping_max;
  double min_relative_cost_reduction, step_factor_decrease, min_step_factor_for_good;
  double absolute_cost_tolerance, relative_cost_tolerance, variables_tolerance;
  int32_t max_step_factor_attempts, try_sub_step;
  int32_t stop_if_no_improvement_for, distance_from_troubled_iteration, max_num_iterations;
  int32_t verbose;
  /* Optimizer::Settings::{solverType, pcgMaxIterations, pcgDesiredResidual} (Optimizer.h:43-45).  solver_type:
   * VB_SOLVER_DIRECT (0) or one of the PCG types (landmark Schur elimination, then preconditioned
   * conjugate gradients on the reduced camera system, Optimizer.cpp:232-331, PCG.cpp:15-97) with the preconditioners of
   * Preconditioner.h: VB_SOLVER_PCG_TRIVIAL (1, identity, :29-47), VB_SOLVER_PCG_JACOBI (2, param-sized diagonal blocks,
   * :52-116), VB_SOLVER_PCG_GAUSS_SEIDEL (3, :118-158: BaSpaCho's pseudo-factor, restated as the Cholesky sweep without
   * Schur updates = symmetric block Gauss-Seidel over the supernode diagonal blocks; BaSpaCho's source is absent, so
   * the block partition and sweep order are this library's), VB_SOLVER_PCG_LOWER_PREC_SOLVE (4, :160-245: complete
   * factor of a float-rounded copy of the reduced system with the reference's diagonal-raise retry; the data and the
   * preconditioned vectors are rounded to float, the arithmetic on them is fp64).  Anything else: VB_ERR_ARG. */
  int32_t solver_type, pcg_max_iterations;
  double pcg_desired_residual;
} vb_settings;
#define VB_SOLVER_DIRECT 0
#define VB_SOLVER_PCG_TRIVIAL 1
#define VB_SOLVER_PCG_JACOBI 2
#define VB_SOLVER_PCG_GAUSS_SEIDEL 3
#define VB_SOLVER_PCG_LOWER_PREC_SOLVE 4

typedef struct vb_summary {  /* Optimizer::Summary, Optimizer.h:93-99 */
  double initial_cost, final_cost;
  int32_t num_troubled_seqs, largest_troubled_seq, num_iterations;
  int32_t stop_reason; /* 0 converged, 1 iteration limit, 2 damping out of range, 3 numeric failure (VB_ERR_NUMERIC; variables left at the current iterate) */
} vb_summary;

typedef struct vb_cost_stats { /* CostStats, Factor.h:25-31 */
  int64_t num_total, num_invalid, num_prev_invalid;
} vb_cost_stats;

typedef struct vb_iter_times { /* ms, CUDA-event timed; the split the reference logs (Optimizer.cpp:1062-1067) */
  double grad_hess, damping, factor_schur, factor_chol, solve, apply, cost, total;
  double step_dot_grad; /* step.dot(grad) before the sign flip (Optimizer.cpp:833): 2 x modelCostReduction */
} vb_iter_times;

/* One LM iteration as the reference logs it (Optimizer.cpp:1058-1076), plus which branches of the loop ran.
 * `damping` is the value AFTER this iteration's adjustment (what the reference prints as "lmbd"),
 * `damping_used` the value the linear system was damped with. */
#define VB_BR_MODEL_NEGATIVE 1     /* Q1: modelCostReduction < 0, g/H recomputed, damping bumped, fell through (:835-855) */
#define VB_BR_STEP_FACTOR 2        /* step-factor loop entered (:907-1011) */
#define VB_BR_ACCEPT_AT_FACTOR 4   /* accepted after scaling the step (:941-957) */
#define VB_BR_SUB_STEP 8           /* sub-step solved on the gradient at the scaled step (:962-975) */
#define VB_BR_ACCEPT_AT_SUB 16     /* accepted after the sub-step (:985-1000) */
#define VB_BR_LATCH_NO_RETRY 32    /* Q5: dontRetryFailed latched in this iteration (:1002-1007) */
#define VB_BR_REJECTED 64          /* newCost > prevCost or failure rate: restore, damping x fail (:1019-1035) */
#define VB_BR_GOOD 128             /* damping x good (:1043-1047) */
#define VB_BR_AVERAGE 256          /* damping x average (:1048-1051) */
#define VB_BR_TOLERANCE_HIT 512    /* one of the three tolerances was hit (:1013-1018) */
#define VB_BR_TROUBLED_SEQ_END 1024 /* a troubled sequence ended and was counted (:1037-1042) */
#define VB_BR_FACTOR_FAILED 2048   /* the damped system was not positive definite in floating point: counted as a rejected step */
typedef struct vb_iter_trace {
  double prev_cost, new_cost, damping, damping_used, model_cost_reduction, applied_step_factor;
  double ratio_reduction_to_expected, cost_reduction, grad_norm, step_norm;
  double ratio_s2v_inf, ratio_s2v_2, ratio_s2v_1;   /* applyStep statistics (Optimizer.cpp:886) */
  double ms_total, ms_grad_hess, ms_factor, ms_solve; /* host wall clock, as the reference measures them */
  int64_t num_total, num_invalid, num_prev_invalid; /* CostStats the decision used (Q3: last store) */
  int32_t branch;                                   /* VB_BR_* bits */
  int32_t step_factor_attempts;
} vb_iter_trace;

/* Hooks of Optimizer::Settings (lib/small_thing/Optimizer.h:81,90): `log` receives the reference's
 * per-iteration text (same layout as Optimizer.cpp:1058-1076 and the convergence / limit / damping lines),
 * `pre_step` runs on the calling thread at the top of every iteration where Optimizer.cpp:803-805 calls
 * settings.preStepCallback.  The device is the source of truth inside vb_optimize (SURVEY.md 8b
 * "Ownership"): before each pre_step call the variables whose kind is in `sync_kinds` (bit k = VB_VAR_k;
 * the reference's callbacks read IMU calibrations and gravity) are copied into `host_vars` (caller-owned,
 * var_data_len doubles, same layout as vb_graph_desc.var_data).  The callback may call vb_upload_store_aux /
 * vb_upload_store_consts / vb_upload_vars / vb_rs_refresh on the same problem; nothing else. */
typedef void (*vb_pre_step_fn)(void* user, int32_t iteration);
typedef void (*vb_log_fn)(void* user, const char* text);
typedef struct vb_optimize_hooks {
  vb_pre_step_fn pre_step; void* pre_step_user;
  uint32_t sync_kinds; double* host_vars;
  vb_log_fn log; void* log_user;
  vb_iter_trace* trace; int32_t trace_cap; /* first trace_cap iterations are recorded */
  int32_t trace_len;                       /* out: iterations recorded */
} vb_optimize_hooks;

typedef struct vb_ctx vb_ctx;
typedef struct vb_problem vb_problem;

int vb_version(void);
const char* vb_last_error(void);
void vb_settings_default(vb_settings* s);

/* device context (one host thread drives one context; Optimizer is not thread-safe either) */
int vb_ctx_create(int device, vb_ctx** out);
/* same, but every kernel and copy is ordered on a caller-owned cudaStream_t (the stream the caller's
 * NCCL communicator is ordered against), so collectives interleave without host synchronisation */
int vb_ctx_create_on_stream(int device, void* cuda_stream, vb_ctx** out);
int vb_ctx_destroy(vb_ctx* ctx);

/* Optimizer::initSolver (Optimizer.cpp:166-207) + FactorStore::registerBlocks (Factor.h:761-785)
 * + BaSpaCho::createSolver: uploads the graph as SoA tables and runs the symbolic phase
 * (block structure, ordering, supernodes, scatter maps).  Host arrays in `g` may be freed after. */
int vb_problem_create(vb_ctx* ctx, const vb_graph_desc* g, vb_problem** out);
int vb_problem_destroy(vb_problem* p);

/* ---- multi-GPU (one process per GPU) ------------------------------------------------
 * Every rank creates the SAME graph with its own (rank, nranks).  The symbolic phase is identical on
 * all ranks.  The elimination tree of the reduced camera system is cut into subtrees dealt to the ranks
 * under a shared top (factored by every rank); each landmark and each factor belongs to the rank whose
 * subtree it touches (vb_shard_plan, vb_shard_factor_owners), which evaluates, eliminates and solves it.  The
 * exchange steps call the hook, which must perform the collective IN PLACE on device memory, ordered
 * on `cuda_stream` (ncclAllReduce(sum, double) / ncclBroadcast(root 0) in a C++ host, torch.distributed
 * in the Python mirror), and return 0:
 *   - factor : sum of the shared top of the reduced camera system, after the rank's own subtrees are factored
 *   - solve  : sums of the camera right-hand side / camera step (each scalar from one rank), sum of the landmark steps
 *   - scalars: cost, CostStats, step.grad (one small sum per host decision)
 * The reference parallelises this path with threads only (Optimizer.cpp:797); there is no
 * reference interface to cite for the exchange. */
#define VB_COLL_SUM 0
#define VB_COLL_BCAST0 1
typedef int (*vb_collective_fn)(void* user, int op, double* dev_ptr, int64_t count, void* cuda_stream);
int vb_problem_create_sharded(vb_ctx* ctx, const vb_graph_desc* g, int rank, int nranks,
                              vb_collective_fn fn, void* user, vb_problem** out);
/* The same with the hook implemented inside the library, in C++, over NCCL (csrc/nccl_hook.cpp): no Python and no torch on
 * the collective path.  NCCL is bound at run time (libnccl.so.2).  Rank 0 makes the 128-byte id (vb_nccl_unique_id) and
 * hands it to the other ranks by whatever means the host has (MPI, a file, torch.distributed in bench.py); every rank
 * then calls vb_nccl_comm_create (collective: ncclCommInitRank) with its CUDA device current, and creates its problem
 * with vb_problem_create_sharded_nccl.  The context must have been created on the stream the caller wants the
 * collectives ordered on (vb_ctx_create_on_stream) or on its own stream (vb_ctx_create). */
int vb_nccl_version(int* version);
int vb_nccl_unique_id(char* id128);
int vb_nccl_comm_create(int rank, int nranks, const char* id128, void** comm_out);
int vb_nccl_comm_destroy(void* comm);
int vb_nccl_comm_stats(void* comm, int64_t* calls, double* bytes);
const char* vb_nccl_last_error(void);
int vb_problem_create_sharded_nccl(vb_ctx* ctx, const vb_graph_desc* g, int rank, int nranks, void* comm, vb_problem** out);
/* host-only: owner rank of every landmark (indexed by registration index, n = num_points) */
int vb_shard_plan(const vb_graph_desc* g, int nranks, int32_t* landmark_owner, int64_t n);
/* host-only: owner rank of every factor of one store (n = its count); *aligned = 1 when no landmark or
 * factor touches the subtrees of two different ranks (then only the shared top of the reduced system is
 * summed over ranks) */
int vb_shard_factor_owners(const vb_graph_desc* g, int nranks, int32_t store, int32_t* factor_owner, int64_t n, int32_t* aligned);
/* out8 = {rank, nranks, owned landmarks, owned visual factors, reduced-camera-system doubles summed per
 * factorisation, collective calls so far, collective bytes so far, lumps of this rank's own subtrees} */
int vb_shard_info(const vb_problem* p, double* out8);

/* Solver::order(), dataSize() (Optimizer.cpp:339-340,777,783-784) and symbolic-phase outputs */
/* host-only symbolic statistics for a given supernode width cap / nested-dissection leaf size:
 * out8 = {order, dataSize, lumps, levels, maxLumpRows, cholFlops, schurFlops, widest level} */
int vb_symbolic_stats(const vb_graph_desc* g, int max_lump_width, int nd_leaf_blocks, double* out8);
/* host-only: padding of the landmark-group Schur update for a grouping: out8 = {groups, tile tasks, executed DMMA
 * flops, algorithmic Schur flops, mean landmarks per group, mean union rows, levels, algorithmic Cholesky flops} */
int vb_schur_plan_stats(const vb_graph_desc* g, int max_rows, int chunks, int nd_leaf_blocks, double* out8);
int64_t vb_order(const vb_problem* p);
int64_t vb_data_size(const vb_problem* p);
int64_t vb_num_blocks(const vb_problem* p);
int64_t vb_num_lumps(const vb_problem* p);
int64_t vb_num_levels(const vb_problem* p);                 /* depth of the supernodal elimination tree */
int64_t vb_num_launches_per_iteration(const vb_problem* p); /* kernels launched by vb_lm_iteration_timed */
/* host-only symbolic phase (no device needed): order, dataSize, lumps, permutation[vb_num_blocks] */
int vb_symbolic_only(const vb_graph_desc* g, int64_t* order, int64_t* data_size, int64_t* num_lumps, int64_t* perm, int64_t perm_len);
/* accessor().paramStart(i) for every registration index (Factor.h:612) */
int vb_get_param_start(const vb_problem* p, int64_t* out, int64_t n);
/* elimination position of every registration index (the permutation) */
int vb_get_perm(const vb_problem* p, int64_t* out, int64_t n);
/* lump (supernode) boundaries in elimination positions: out[num_lumps+1] */
int vb_get_lump_start(const vb_problem* p, int64_t* out, int64_t n);
/* accessor().block(data,i,j): offset + row stride of block (i,j) (registration indices, any order);
 * *transposed = 1 when the stored block is (j,i).  offset < 0 if the block is not in the structure. */
int vb_block_offset(const vb_problem* p, int64_t i, int64_t j, int64_t* offset, int64_t* stride, int32_t* transposed);
/* flops / bytes of the numeric phases from the symbolic analysis */
int vb_get_work(const vb_problem* p, double* schur_flops, double* chol_flops, double* visual_bytes);

/* Optimizer::computeGradHess (Optimizer.cpp:57-71 -> Factor.h:704-734,543-661).
 * Zeroes grad (and hess when with_hess) first, as Optimizer.cpp:807-808 does. */
int vb_eval_grad_hess(vb_problem* p, int with_hess, int update_cache, int dont_retry_failed, double* cost);
/* Optimizer::computeCost (Optimizer.cpp:88-97 -> Factor.h:664-701,390-417) */
int vb_eval_cost(vb_problem* p, int comparable, double* cost, vb_cost_stats* stats);
/* Optimizer::addDamping (Optimizer.cpp:136-146) */
int vb_damp(vb_problem* p, double lambda);
/* step = grad (Optimizer.cpp:828) */
int vb_step_from_grad(vb_problem* p);
/* Solver::factor(hess) (Optimizer.cpp:223): landmark Schur elimination + supernodal Cholesky */
int vb_factor(vb_problem* p);
/* Solver::solve(hess, vec, order, 1) on the step vector (Optimizer.cpp:228); which=0 step, 1 grad_new */
int vb_solve(vb_problem* p, int which);
/* step.dot(grad), |grad|, |step| (Optimizer.cpp:833,858-859) */
int vb_step_stats(vb_problem* p, double* step_dot_grad, double* grad_norm, double* step_norm);
/* vec *= s (Optimizer.cpp:857,925,971); which=0 step, 1 grad_new */
int vb_scale_vec(vb_problem* p, int which, double s);
/* -grad_new . step * 0.5 support (Optimizer.cpp:921) */
int vb_dot(vb_problem* p, int a, int b, double* out);
/* Optimizer::applyStep (Optimizer.cpp:121-134 -> Variable.h:352-370): ratios = {inf, rms, mean} */
int vb_apply_step(vb_problem* p, int which, double ratios[3]);
/* Optimizer::backupVariables / restoreVariables (Optimizer.cpp:107-119) */
int vb_backup(vb_problem* p);
int vb_restore(vb_problem* p);
/* compute grad only into the secondary vector (Optimizer.cpp:912-919) */
int vb_eval_grad_new(vb_problem* p, int dont_retry_failed);

/* Linear solver of vb_factor / vb_solve (vb_optimize sets it from its settings): see vb_settings.solver_type.  With a PCG
 * type vb_factor eliminates the landmarks and initialises the preconditioner (Optimizer.cpp:289-300), vb_solve runs the
 * landmark forward solve, PCG on the reduced system and the landmark back-substitution (:301-327); vb_pcg_last returns
 * the iteration count and relative residual of the last PCG run.  Not available on a sharded problem. */
int vb_set_linear_solver(vb_problem* p, int32_t solver_type, int32_t pcg_max_iterations, double pcg_desired_residual);
int vb_pcg_last(const vb_problem* p, int32_t* iterations, double* relative_residual);
/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  *out_dev = a . b over two
 * caller-owned DEVICE vectors of n doubles, launched on `cuda_stream` (a cudaStream_t; NULL = the legacy default stream)
 * of GPU `device`; returns without synchronising.  The allocation-free variant of the dot-product kernel behind vb_dot:
 * its scratch (one partial sum per 16 384 elements + a ticket) is a fixed 128.25 KiB __device__ array that every GPU
 * loading the library holds until exit, so the call performs no cudaMalloc / cudaFree and no allocating library call.
 * n > VB_DOT_NO_ALLOC_MAX_N (or n < 0, or a null pointer) is VB_ERR_ARG with a text saying so; n = 0 gives 0.
 * The result is deterministic (fixed summation order).  All calls on one GPU share the scratch: calls whose kernels
 * OVERLAP on the GPU (different streams without an ordering between them) race; calls ordered on one stream or by events
 * are safe.  The first call on a device looks the array's address up (cudaGetSymbolAddress): make one call before
 * starting a graph capture. */
#define VB_DOT_NO_ALLOC_MAX_N ((int64_t)268435456) /* 16 384 partial sums x 16 384 elements each = 2^28 */
int vb_dot_no_alloc(int device, const double* a_dev, const double* b_dev, int64_t n, double* out_dev, void* cuda_stream);
/* Optimizer::optimize (Optimizer.cpp:768-1106): the whole LM loop, host-controlled, device-resident */
int vb_optimize(vb_problem* p, const vb_settings* s, vb_summary* out);
/* same loop with the Settings hooks: preStepCallback (Optimizer.h:90, called at Optimizer.cpp:803-805),
 * log (Optimizer.h:81) and a per-iteration trace; hooks may be NULL */
int vb_optimize_ex(vb_problem* p, const vb_settings* s, vb_optimize_hooks* hooks, vb_summary* out);
/* one accepted-path LM iteration with per-phase CUDA-event timing (bench / profiling) */
int vb_lm_iteration_timed(vb_problem* p, double lambda, int apply, vb_iter_times* t, double* cost_before, double* cost_after);

/* data movement.  On a sharded problem the variables and the step are complete and bit-identical on every
 * rank; grad holds this rank's contributions (vb_step_stats sums it on demand) and hess this rank's
 * panels: its own landmarks and subtrees, plus the summed shared top after vb_factor. */
int vb_download_vars(vb_problem* p, double* var_data, int64_t n);       /* all variable data */
int vb_upload_vars(vb_problem* p, const double* var_data, int64_t n);
int vb_download_grad(vb_problem* p, double* out, int64_t n);             /* permuted order */
int vb_download_step(vb_problem* p, double* out, int64_t n);
int vb_download_hess(vb_problem* p, double* out, int64_t n);
int vb_upload_hess(vb_problem* p, const double* in, int64_t n);
int vb_upload_grad(vb_problem* p, const double* in, int64_t n);
int vb_download_factor_cache(vb_problem* p, int32_t store, double* out, int64_t n); /* ResultCache, Factor.h:92-94 */
int vb_upload_store_aux(vb_problem* p, int32_t store, const double* aux, int64_t n); /* refreshed preint / RS tables */
int vb_upload_store_consts(vb_problem* p, int32_t store, const double* consts, int64_t n);
int vb_download_store_aux(vb_problem* p, int32_t store, double* out, int64_t n);

/* Per-iteration refresher of the rolling-shutter tables on the device:
 * RollingShutterData::compute (lib/motion/preintegration/RollingShutterData.cpp:16-65), which the reference
 * runs on the host for every RS rig at the top of every LM iteration
 * (SingleSessionAdapter::updateRollingShutterData, viba/single_session/InitCalibration.cpp:316-325, from the
 * preStepCallback at Optimizer.cpp:801-805).
 * vb_rs_set_imu hands the raw IMU-0 samples (time in ns, float rad/s and m/s^2 as in ImuMeasurement) and,
 * per table row of a VB_F_VISUAL_RS store, the rig's timestamp in us and the pool id of the IMU calibration
 * variable of its window (-1: leave that row alone); half_length_us is intervalHalfLengthUs_ (readout / 2 +
 * 2 ms in the reference).  From then on vb_optimize recomputes the tables from the current calibration at the
 * top of every iteration (vb_lm_iteration_timed too, inside its grad_hess phase); vb_rs_refresh does it on request;
 * n_samples = 0 forgets the samples of that store.  VB_ERR_NUMERIC when a window runs off the
 * samples, holds more than 32 gyro boundaries, or sample times do not increase (the reference throws). */
int vb_rs_set_imu(vb_problem* p, int32_t store, int64_t n_rigs, const int64_t* rig_mid_time_us, const int32_t* rig_calib_var,
                  int64_t half_length_us, int64_t n_samples, const int64_t* sample_time_ns, const float* gyro_rad_s,
                  const float* accel_m_s2);
int vb_rs_refresh(vb_problem* p, int32_t store); /* store < 0: every store that has IMU samples */
/* updateRollingShutterData also re-snapshots gravityWorld from the gravity variable (InitCalibration.cpp:317 ->
 * RollingShutterData.cpp:21): name that S2 variable here and every refresh of the store copies its current value into the
 * store's aux[0..2]; -1 keeps the value uploaded with the store (single-session runs hold gravity constant) */
int vb_rs_set_gravity_var(vb_problem* p, int32_t store, int32_t gravity_var);

/* Optimizer::computeJointCovariances / computeCovariances (lib/small_thing/Optimizer.cpp:503-696; used by
 * SingleSessionProblem::computeCovariances, viba/problem/SingleSessionProblem.cpp:66-138) for ONE list of registered
 * variables (registration indices): cov = the block of (H + damping)^-1 of those variables, row-major D x D with
 * D = sum of their tangent dims in list order (offsets[n + 1], may be null); damping follows Settings::damping
 * (1e-5 by default in the reference) and is raised as the reference does when the factorisation fails (*damping_used,
 * may be null).  Linearises at the current values without touching the ResultCache; overwrites grad / hess / step. */
int vb_compute_joint_covariance(vb_problem* p, double damping, int64_t n, const int64_t* indices, double* cov, int64_t cov_len,
                                int64_t* offsets, double* damping_used);

/* The same for MANY lists in one call, as SingleSessionProblem::computeCovariances uses it (SingleSessionProblem.cpp:100-126:
 * one list per rig {pose, velocity[, omega]} + one per calibration variable): list l = indices[list_ptr[l] .. list_ptr[l + 1]);
 * the system is linearised, damped and factored ONCE; cov receives the D_l x D_l blocks back to back (cov_ptr[n_lists + 1],
 * may be null, gives where each starts). */
int vb_compute_joint_covariances(vb_problem* p, double damping, int64_t n_lists, const int64_t* list_ptr, const int64_t* indices, double* cov,
                                 int64_t cov_len, int64_t* cov_ptr, double* damping_used);

/* Optimizer::computeMarginalProblem (Optimizer.cpp:422-494, the quantity TestOptimizer.cpp:74-82 multiplies with the joint
 * covariance): the marginal problem of ONE list of registered variables — H = Schur complement of every other variable
 * onto the listed ones (row-major D x D, blocks in list order, offsets[n + 1] may be null), b = the matching gradient,
 * *cost = the cost minus what the eliminated variables can still remove (0.5 g_r^T H_rr^-1 g_r).  Damping and its retry as
 * in vb_compute_joint_covariance; same side effects.  The reference reorders the listed variables last and reads a
 * partial factorisation; here the ordering is fixed at problem creation and the same quantities come from complete
 * solves (H = inverse of the joint covariance block, inverted on the host: D is small). */
int vb_compute_marginal_problem(vb_problem* p, double damping, int64_t n, const int64_t* indices, double* H, int64_t h_len, double* b,
                                int64_t b_len, double* cost, int64_t* offsets, double* damping_used);

/* refinePoints pre-pass: viba/problem/PointRefinement.cpp:96-196 — a 3x3 Levenberg-Marquardt per
 * landmark (lambda 1e-5, <= 5 iterations, reduced step on a poor one) over its visual factors with every other variable
 * held fixed (FactorStore::varGradHess / singleCost, lib/small_thing/Factor.h:230-232,419-468).  Updates the landmark
 * variables on the device; `out` receives the totals the reference logs (:180-193); per_point (may be null) receives
 * {startCost, endCost, nIts (-1: failed)} for each of the vb_refine_num_points landmarks that have visual factors, in
 * ascending variable id.  Not available on a sharded problem. */
typedef struct vb_refine_stats {
  double total_start_cost, total_end_cost;
  int64_t num_points, num_failures, num_total_iterations, num_at_least_one_iteration;
} vb_refine_stats;
int vb_refine_points(vb_problem* p, vb_refine_stats* out, double* per_point, int64_t n_points);
int64_t vb_refine_num_points(const vb_problem* p);
/* Triangulation refinement (second half): SingleSessionAdapter::refineTriangulationResult,
 * viba/single_session/Triangulation.cpp:99-161, as triangulatePoint runs it twice (:219-238; settings of pass 1 / 2 in
 * Triangulation.h:29-39: 3.0, false, 3, 1.5 / 2.5, true, 3, 1.0) — for every landmark with visual factors at once:
 * max_num_iterations undamped Gauss-Newton steps on the point with HuberLoss(huber_radius) on the weighted reprojection
 * error, re-classifying the observations every iteration (inlier = unweighted image error below outlier_threshold pixels)
 * and, with skip_outliers, leaving the outliers out of the normal equations.  Updates the landmark variables on the
 * device.  per_point (may be null): {numInlierObservations, cost of the last iteration, status} per landmark, same order
 * and count as vb_refine_points; status -1 = the 3x3 system was not positive definite (Eigen's llt() returns garbage
 * there; here the point is kept), counted in *num_failures (may be null).  vb_refine_triangulation_inliers gives the
 * inlierObservationIndices of the last call as one byte per factor of a visual store, in factor order.  The RANSAC
 * candidate search of triangulatePoint (:33-97) and the track bookkeeping stay on the host (session builder, not
 * restated).  Not available on a sharded problem. */
int vb_refine_triangulation(vb_problem* p, double outlier_threshold, int32_t skip_outliers, int32_t max_num_iterations, double huber_radius,
                            double* per_point, int64_t n_points, int64_t* num_failures);
int vb_refine_triangulation_inliers(vb_problem* p, int32_t store, uint8_t* flags, int64_t n_factors);

/* Per-iteration refresher of the preintegration records on the device (second half):
 * computePreIntegration (lib/motion/preintegration/PreIntegration.cpp:136-275, with CompensateJac.cpp:146-249 and
 * MotionIntegral.cpp:52-75,162-226), which the reference runs on the host for every inertial factor at the top of
 * every LM iteration under --recompute-preint (SingleSessionAdapter::regenerateAllPreintegrationsFromImuMeasurements,
 * viba/single_session/InertialFactors.cpp:19-70, from the preStepCallback of interfaces/ark/main_AriaKit_ViBa.cpp:95-101).
 * vb_preint_set_imu hands the raw samples of the store's IMU and, per factor of a VB_F_INERTIAL* store, the interval
 * [start_us, end_us] (previous / next rig timestamps) and the pool id of the IMU calibration variable whose current
 * value is the measurement model (-1: leave that record alone); gyro_var3 / accel_var3 are
 * ImuNoiseModelParameters::{gyroSampleVarianceRad2Sec2, accelSampleVarianceM2Sec4}.  From then on vb_optimize (and
 * vb_lm_iteration_timed, inside its grad_hess phase) rewrites the store's records {R, dV, dP, dt, J(9x23), evalPoint}
 * before the rolling-shutter tables, as the reference's callback does; vb_preint_refresh does it on request;
 * vb_preint_download returns the 9x9 covariances (whose inverse is the factor's information matrix at creation,
 * viba/problem/InertialFactor.cpp:313 - NOT refreshed, as in the reference) and omegaAtEnd of the last refresh.
 * n_samples = 0 forgets the samples of that store.  VB_ERR_NUMERIC when an interval runs off the samples. */
int vb_preint_set_imu(vb_problem* p, int32_t store, int64_t n_factors, const int64_t* start_us, const int64_t* end_us,
                      const int32_t* calib_var, int64_t n_samples, const int64_t* sample_time_ns, const float* gyro_rad_s,
                      const float* accel_m_s2, const double* gyro_var3, const double* accel_var3);
int vb_preint_refresh(vb_problem* p, int32_t store); /* store < 0: every store that has IMU samples */
int vb_preint_download(vb_problem* p, int32_t store, double* cov81_per_factor, double* omega_at_end3_per_factor, int64_t n_factors);

/* FP64 DMMA peak probe (mma.sync m8n8k4 f64) used as the tensor roofline denominator */
int vb_measure_dmma_peak(vb_ctx* ctx, double* tflops);
/* the same probe for every FP64 mma.sync shape: tflops4 = {m8n8k4, m16n8k4, m16n8k8, m16n8k16} */
int vb_measure_dmma_shapes(vb_ctx* ctx, double* tflops4);

#ifdef __cplusplus
}
#endif
#endif /* VIBA_B200_H_ */
