"""ctypes binding of libviba_b200.so — the C ABI declared in include/viba_b200.h.

This is the same stub a reference maintainer would bind; it holds no
arithmetic.  Loading fails loudly if the CUDA library has not been built: there is no fallback.
"""
import ctypes as C
import os
import numpy as np
from . import graph as G

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libviba_b200.so")
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int64)

EXPORTS = [
    "vb_version", "vb_last_error", "vb_settings_default", "vb_ctx_create", "vb_ctx_destroy",
    "vb_problem_create", "vb_problem_destroy", "vb_order", "vb_data_size", "vb_num_blocks", "vb_num_lumps",
    "vb_num_levels", "vb_num_launches_per_iteration", "vb_symbolic_only", "vb_symbolic_stats",
    "vb_get_param_start", "vb_get_perm", "vb_get_lump_start", "vb_block_offset", "vb_get_work",
    "vb_eval_grad_hess", "vb_eval_cost", "vb_damp", "vb_step_from_grad", "vb_factor", "vb_solve",
    "vb_step_stats", "vb_scale_vec", "vb_dot", "vb_apply_step", "vb_backup", "vb_restore", "vb_eval_grad_new",
    "vb_optimize", "vb_optimize_ex", "vb_lm_iteration_timed", "vb_download_vars", "vb_upload_vars", "vb_download_grad",
    "vb_download_step", "vb_download_hess", "vb_upload_hess", "vb_upload_grad", "vb_download_factor_cache",
    "vb_upload_store_aux", "vb_upload_store_consts", "vb_measure_dmma_peak",
    "vb_download_store_aux", "vb_rs_set_imu", "vb_rs_refresh",
    "vb_dot_no_alloc", "vb_schur_plan_stats", "vb_measure_dmma_shapes", "vb_ctx_create_on_stream", "vb_problem_create_sharded", "vb_shard_plan", "vb_shard_info", "vb_shard_factor_owners",
]
# int fn(void* user, int op, double* dev_ptr, int64_t count, void* cuda_stream)  (include/viba_b200.h)
COLLECTIVE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p)
COLL_SUM, COLL_BCAST0 = 0, 1

_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(make -C visual_inertial_bundle_adjustment_b200/csrc). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    lib.vb_last_error.restype = C.c_char_p
    for f in ("vb_order", "vb_data_size", "vb_num_blocks", "vb_num_lumps", "vb_num_levels", "vb_num_launches_per_iteration"):
        getattr(lib, f).restype = C.c_int64
        getattr(lib, f).argtypes = [C.c_void_p]
    lib.vb_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    lib.vb_ctx_destroy.argtypes = [C.c_void_p]
    lib.vb_ctx_create_on_stream.argtypes = [C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]
    lib.vb_problem_create_sharded.argtypes = [C.c_void_p, C.POINTER(G.vb_graph_desc), C.c_int, C.c_int, COLLECTIVE_FN,
                                              C.c_void_p, C.POINTER(C.c_void_p)]
    lib.vb_shard_plan.argtypes = [C.POINTER(G.vb_graph_desc), C.c_int, C.POINTER(C.c_int32), C.c_int64]
    lib.vb_shard_info.argtypes = [C.c_void_p, _dp]
    lib.vb_shard_factor_owners.argtypes = [C.POINTER(G.vb_graph_desc), C.c_int, C.c_int32, C.POINTER(C.c_int32), C.c_int64,
                                           C.POINTER(C.c_int32)]
    lib.vb_problem_create.argtypes = [C.c_void_p, C.POINTER(G.vb_graph_desc), C.POINTER(C.c_void_p)]
    lib.vb_problem_destroy.argtypes = [C.c_void_p]
    lib.vb_get_param_start.argtypes = [C.c_void_p, _ip, C.c_int64]
    lib.vb_get_perm.argtypes = [C.c_void_p, _ip, C.c_int64]
    lib.vb_get_lump_start.argtypes = [C.c_void_p, _ip, C.c_int64]
    lib.vb_block_offset.argtypes = [C.c_void_p, C.c_int64, C.c_int64, _ip, _ip, C.POINTER(C.c_int32)]
    lib.vb_get_work.argtypes = [C.c_void_p, _dp, _dp, _dp]
    lib.vb_eval_grad_hess.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, _dp]
    lib.vb_eval_cost.argtypes = [C.c_void_p, C.c_int, _dp, C.POINTER(G.vb_cost_stats)]
    lib.vb_damp.argtypes = [C.c_void_p, C.c_double]
    lib.vb_step_from_grad.argtypes = [C.c_void_p]
    lib.vb_factor.argtypes = [C.c_void_p]
    lib.vb_dot_no_alloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    lib.vb_set_linear_solver.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_double]
    lib.vb_pcg_last.argtypes = [C.c_void_p, C.POINTER(C.c_int32), _dp]
    lib.vb_solve.argtypes = [C.c_void_p, C.c_int]
    lib.vb_step_stats.argtypes = [C.c_void_p, _dp, _dp, _dp]
    lib.vb_scale_vec.argtypes = [C.c_void_p, C.c_int, C.c_double]
    lib.vb_dot.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp]
    lib.vb_apply_step.argtypes = [C.c_void_p, C.c_int, _dp]
    lib.vb_backup.argtypes = [C.c_void_p]
    lib.vb_restore.argtypes = [C.c_void_p]
    lib.vb_eval_grad_new.argtypes = [C.c_void_p, C.c_int]
    lib.vb_optimize.argtypes = [C.c_void_p, C.POINTER(G.vb_settings), C.POINTER(G.vb_summary)]
    lib.vb_optimize_ex.argtypes = [C.c_void_p, C.POINTER(G.vb_settings), C.POINTER(G.vb_optimize_hooks), C.POINTER(G.vb_summary)]
    lib.vb_lm_iteration_timed.argtypes = [C.c_void_p, C.c_double, C.c_int, C.POINTER(G.vb_iter_times), _dp, _dp]
    for f in ("vb_download_vars", "vb_upload_vars", "vb_download_grad", "vb_download_step", "vb_download_hess",
              "vb_upload_hess", "vb_upload_grad"):
        getattr(lib, f).argtypes = [C.c_void_p, _dp, C.c_int64]
    for f in ("vb_download_factor_cache", "vb_upload_store_aux", "vb_upload_store_consts", "vb_download_store_aux"):
        getattr(lib, f).argtypes = [C.c_void_p, C.c_int32, _dp, C.c_int64]
    lib.vb_measure_dmma_peak.argtypes = [C.c_void_p, _dp]
    lib.vb_rs_set_imu.argtypes = [C.c_void_p, C.c_int32, C.c_int64, _ip, C.POINTER(C.c_int32), C.c_int64, C.c_int64, _ip,
                                  C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.vb_rs_refresh.argtypes = [C.c_void_p, C.c_int32]
    lib.vb_symbolic_only.argtypes = [C.POINTER(G.vb_graph_desc), _ip, _ip, _ip, _ip, C.c_int64]
    lib.vb_schur_plan_stats.argtypes = [C.POINTER(G.vb_graph_desc), C.c_int, C.c_int, C.c_int, _dp]
    lib.vb_symbolic_stats.argtypes = [C.POINTER(G.vb_graph_desc), C.c_int, C.c_int, _dp]
    _lib = lib
    return lib


class VbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"viba_b200 error {code}: {msg}")
        self.code = code


def _ck(lib, rc):
    if rc != 0:
        raise VbError(rc, lib.vb_last_error().decode())


def _d(a):
    return a.ctypes.data_as(_dp)


def symbolic_only(graph):
    """Host-only symbolic phase (no GPU needed): order, dataSize, number of lumps, permutation."""
    lib = load()
    desc = graph.desc()
    o, ds, nl = C.c_int64(), C.c_int64(), C.c_int64()
    perm = np.zeros(graph.num_blocks, np.int64)
    _ck(lib, lib.vb_symbolic_only(C.byref(desc), C.byref(o), C.byref(ds), C.byref(nl), perm.ctypes.data_as(_ip), perm.size))
    return o.value, ds.value, nl.value, perm


DOT_NO_ALLOC_MAX_N = 1 << 28


def dot_no_alloc(device, a_ptr, b_ptr, n, out_ptr, stream=None):
    """vb_dot_no_alloc: *out_ptr = a . b over raw DEVICE pointers (ints), on `stream` (raw cudaStream_t or None), without
    any allocation and without synchronising; n <= DOT_NO_ALLOC_MAX_N"""
    lib = load()
    _ck(lib, lib.vb_dot_no_alloc(int(device), C.c_void_p(a_ptr), C.c_void_p(b_ptr), int(n), C.c_void_p(out_ptr), C.c_void_p(stream or 0)))


def symbolic_stats(graph, max_lump_width=0, nd_leaf_blocks=600):
    lib = load()
    desc = graph.desc()
    out = np.zeros(8)
    _ck(lib, lib.vb_symbolic_stats(C.byref(desc), max_lump_width, nd_leaf_blocks, _d(out)))
    return dict(zip(["order", "data_size", "lumps", "levels", "max_lump_rows", "chol_flops", "schur_flops", "widest_level"], out))


def schur_plan_stats(graph, max_rows=1024, chunks=4, nd_leaf_blocks=0):
    """Host-only: executed vs algorithmic flops of the landmark-group Schur update for a grouping."""
    lib = load()
    desc = graph.desc()
    out = np.zeros(8)
    _ck(lib, lib.vb_schur_plan_stats(C.byref(desc), max_rows, chunks, nd_leaf_blocks, _d(out)))
    return dict(zip(["groups", "tasks", "executed_flops", "schur_flops", "landmarks_per_group", "union_rows", "levels", "chol_flops"], out))


def shard_plan(graph, nranks):
    """Host-only: owner rank of every landmark (registration index) for an nranks-way split."""
    lib = load()
    desc = graph.desc()
    out = np.zeros(graph.num_points, np.int32)
    _ck(lib, lib.vb_shard_plan(C.byref(desc), nranks, out.ctypes.data_as(C.POINTER(C.c_int32)), out.size))
    return out


def shard_factor_owners(graph, nranks, store):
    """Host-only: (owner rank of every factor of `store`, plan aligned with the elimination tree?)."""
    lib = load()
    desc = graph.desc()
    out = np.zeros(graph.stores[store].count, np.int32)
    al = C.c_int32()
    _ck(lib, lib.vb_shard_factor_owners(C.byref(desc), nranks, store, out.ctypes.data_as(C.POINTER(C.c_int32)), out.size, C.byref(al)))
    return out, bool(al.value)


class Context:
    def __init__(self, device=0, stream=None):
        """`stream`: a raw cudaStream_t handle (int) every kernel is ordered on (sharded runs pass
        the stream their collectives are ordered against); default is a private stream."""
        self.lib = load()
        self.h = C.c_void_p()
        self.device = device
        if stream is None:
            _ck(self.lib, self.lib.vb_ctx_create(device, C.byref(self.h)))
        else:
            _ck(self.lib, self.lib.vb_ctx_create_on_stream(device, C.c_void_p(stream), C.byref(self.h)))

    def dmma_peak_tflops(self):
        t = C.c_double()
        _ck(self.lib, self.lib.vb_measure_dmma_peak(self.h, C.byref(t)))
        return t.value

    def dmma_shapes_tflops(self):
        out = np.zeros(4)
        self.lib.vb_measure_dmma_shapes.argtypes = [C.c_void_p, _dp]
        _ck(self.lib, self.lib.vb_measure_dmma_shapes(self.h, _d(out)))
        return dict(zip(["m8n8k4", "m16n8k4", "m16n8k8", "m16n8k16"], out.tolist()))

    def close(self):
        if self.h:
            self.lib.vb_ctx_destroy(self.h)
            self.h = C.c_void_p()


def nccl_unique_id():
    """128-byte NCCL id for vb_nccl_comm_create (rank 0 makes it, every rank needs it)"""
    lib = load()
    buf = C.create_string_buffer(128)
    if lib.vb_nccl_unique_id(buf) != 0:
        lib.vb_nccl_last_error.restype = C.c_char_p
        raise RuntimeError("NCCL: " + lib.vb_nccl_last_error().decode())
    return buf.raw


class NcclComm:
    """an NCCL communicator owned by the library's C++ hook (csrc/nccl_hook.cpp); creation is collective"""

    def __init__(self, rank, nranks, id128):
        self.lib = load()
        self.h = C.c_void_p()
        self.lib.vb_nccl_comm_create.argtypes = [C.c_int, C.c_int, C.c_char_p, C.POINTER(C.c_void_p)]
        if self.lib.vb_nccl_comm_create(rank, nranks, id128, C.byref(self.h)) != 0:
            self.lib.vb_nccl_last_error.restype = C.c_char_p
            raise RuntimeError("NCCL: " + self.lib.vb_nccl_last_error().decode())

    @property
    def stats(self):
        calls, nbytes = C.c_int64(0), C.c_double(0)
        self.lib.vb_nccl_comm_stats.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_double)]
        self.lib.vb_nccl_comm_stats(self.h, C.byref(calls), C.byref(nbytes))
        return calls.value, nbytes.value

    calls = property(lambda s: s.stats[0])
    bytes = property(lambda s: s.stats[1])

    def close(self):
        if self.h:
            self.lib.vb_nccl_comm_destroy.argtypes = [C.c_void_p]
            self.lib.vb_nccl_comm_destroy(self.h)
            self.h = C.c_void_p()


class Problem:
    """Thin object wrapper over vb_problem; method names follow small_thing::Optimizer
    (lib/small_thing/Optimizer.h:115-335) where one exists."""

    def __init__(self, ctx, graph, rank=0, nranks=1, collective=None, nccl_comm=None):
        """`collective(op, dev_ptr, count, stream) -> None` performs the in-place exchange of a sharded
        problem (sharding.TorchCollective for torch.distributed/NCCL); `nccl_comm` (a capi.NcclComm) selects the
        library's own C++ NCCL hook instead: no Python on the collective path."""
        self.ctx, self.lib, self.g = ctx, ctx.lib, graph
        self.h = C.c_void_p()
        desc = graph.desc()
        self.rank, self.nranks = rank, nranks
        if nranks == 1:
            _ck(self.lib, self.lib.vb_problem_create(ctx.h, C.byref(desc), C.byref(self.h)))
        elif nccl_comm is not None:
            self._comm = nccl_comm
            self.lib.vb_problem_create_sharded_nccl.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]
            _ck(self.lib, self.lib.vb_problem_create_sharded_nccl(ctx.h, C.byref(desc), rank, nranks, nccl_comm.h, C.byref(self.h)))
        else:
            self._coll_error = None

            def hook(user, op, ptr, count, stream):
                try:
                    collective(op, ptr, count, stream)
                    return 0
                except BaseException as ex:  # never let an exception cross the C ABI
                    self._coll_error = ex
                    return 1
            self._hook = COLLECTIVE_FN(hook)  # keep alive as long as the problem
            _ck(self.lib, self.lib.vb_problem_create_sharded(ctx.h, C.byref(desc), rank, nranks, self._hook, None, C.byref(self.h)))
        self.order = self.lib.vb_order(self.h)
        self.data_size = self.lib.vb_data_size(self.h)
        self.num_blocks = self.lib.vb_num_blocks(self.h)
        self.num_lumps = self.lib.vb_num_lumps(self.h)
        self.num_levels = self.lib.vb_num_levels(self.h)
        self.launches_per_iteration = self.lib.vb_num_launches_per_iteration(self.h)

    def close(self):
        if self.h:
            self.lib.vb_problem_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def param_start(self):
        out = np.zeros(self.num_blocks, np.int64)
        _ck(self.lib, self.lib.vb_get_param_start(self.h, out.ctypes.data_as(_ip), out.size))
        return out

    def perm(self):
        out = np.zeros(self.num_blocks, np.int64)
        _ck(self.lib, self.lib.vb_get_perm(self.h, out.ctypes.data_as(_ip), out.size))
        return out

    def lump_start(self):
        out = np.zeros(self.num_lumps + 1, np.int64)
        _ck(self.lib, self.lib.vb_get_lump_start(self.h, out.ctypes.data_as(_ip), out.size))
        return out

    def block_offset(self, i, j):
        o, s, t = C.c_int64(), C.c_int64(), C.c_int32()
        _ck(self.lib, self.lib.vb_block_offset(self.h, i, j, C.byref(o), C.byref(s), C.byref(t)))
        return o.value, s.value, t.value

    def shard_info(self):
        out = np.zeros(8)
        _ck(self.lib, self.lib.vb_shard_info(self.h, _d(out)))
        return dict(zip(["rank", "nranks", "owned_landmarks", "owned_visual", "reduced_doubles", "coll_calls",
                         "coll_bytes", "own_lumps"], out))

    def work(self):
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        _ck(self.lib, self.lib.vb_get_work(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return dict(schur_flops=a.value, chol_flops=b.value, visual_bytes=c.value)

    def compute_grad_hess(self, with_hess=True, update_cache=True, dont_retry_failed=False):
        c = C.c_double()
        _ck(self.lib, self.lib.vb_eval_grad_hess(self.h, int(with_hess), int(update_cache), int(dont_retry_failed), C.byref(c)))
        return c.value

    def compute_grad_new(self, dont_retry_failed=False):
        _ck(self.lib, self.lib.vb_eval_grad_new(self.h, int(dont_retry_failed)))

    def compute_cost(self, comparable=True):
        c, st = C.c_double(), G.vb_cost_stats()
        _ck(self.lib, self.lib.vb_eval_cost(self.h, int(comparable), C.byref(c), C.byref(st)))
        return c.value, (st.num_total, st.num_invalid, st.num_prev_invalid)

    def add_damping(self, lam): _ck(self.lib, self.lib.vb_damp(self.h, lam))
    def step_from_grad(self): _ck(self.lib, self.lib.vb_step_from_grad(self.h))
    def factor(self): _ck(self.lib, self.lib.vb_factor(self.h))
    def solve(self, which=0): _ck(self.lib, self.lib.vb_solve(self.h, which))
    def scale(self, which, s): _ck(self.lib, self.lib.vb_scale_vec(self.h, which, s))
    def backup_variables(self): _ck(self.lib, self.lib.vb_backup(self.h))
    def restore_variables(self): _ck(self.lib, self.lib.vb_restore(self.h))

    def dot(self, a, b):
        o = C.c_double()
        _ck(self.lib, self.lib.vb_dot(self.h, a, b, C.byref(o)))
        return o.value

    def apply_step(self, which=0):
        r = np.zeros(3)
        _ck(self.lib, self.lib.vb_apply_step(self.h, which, _d(r)))
        return r

    def optimize(self, settings):
        s = G.vb_summary()
        _ck(self.lib, self.lib.vb_optimize(self.h, C.byref(settings), C.byref(s)))
        return s

    def optimize_ex(self, settings, trace_cap=300, pre_step=None, sync_kinds=0, log=None):
        """Optimizer::optimize with the Settings hooks (vb_optimize_ex): `pre_step(iteration, host_vars)` runs at
        the top of every iteration with the variables of the kinds in `sync_kinds` (bit k = VB_VAR_k) fresh in
        `host_vars`; `log(text)` receives the reference's per-iteration text.  Returns (summary, trace records)."""
        s = G.vb_summary()
        h = G.vb_optimize_hooks()
        tr = (G.vb_iter_trace * max(1, trace_cap))()
        h.trace, h.trace_cap = tr, trace_cap
        host = np.array(self.g.var_data, dtype=np.float64, copy=True)
        err = []
        if pre_step is not None:
            def _pre(user, it):
                try:
                    pre_step(it, host)
                except BaseException as ex:   # never let an exception cross the C ABI
                    err.append(ex)
            h.pre_step = G.PRE_STEP_FN(_pre)
            h.sync_kinds, h.host_vars = sync_kinds, _d(host)
        if log is not None:
            h.log = G.LOG_FN(lambda user, text: log(text.decode("utf-8", "replace")))
        rc = self.lib.vb_optimize_ex(self.h, C.byref(settings), C.byref(h), C.byref(s))
        if err:
            raise err[0]
        _ck(self.lib, rc)
        return s, G.trace_records(tr, h.trace_len)

    def lm_iteration_timed(self, lam, apply=False):
        t, c0, c1 = G.vb_iter_times(), C.c_double(), C.c_double()
        _ck(self.lib, self.lib.vb_lm_iteration_timed(self.h, lam, int(apply), C.byref(t), C.byref(c0), C.byref(c1)))
        return {n: getattr(t, n) for n, _ in G.vb_iter_times._fields_}, c0.value, c1.value

    def _get(self, fn, n):
        out = np.zeros(n)
        _ck(self.lib, fn(self.h, _d(out), n))
        return out

    def _put(self, fn, a):
        a = np.ascontiguousarray(a, np.float64)
        _ck(self.lib, fn(self.h, _d(a), a.size))

    vars = property(lambda s: s._get(s.lib.vb_download_vars, s.g.var_data.size))
    grad = property(lambda s: s._get(s.lib.vb_download_grad, s.order))
    step = property(lambda s: s._get(s.lib.vb_download_step, s.order))
    hess = property(lambda s: s._get(s.lib.vb_download_hess, s.data_size))
    def set_vars(self, a): self._put(self.lib.vb_upload_vars, a)
    def set_hess(self, a): self._put(self.lib.vb_upload_hess, a)
    def set_grad(self, a): self._put(self.lib.vb_upload_grad, a)

    def factor_cache(self, store):
        out = np.zeros(self.g.stores[store].count)
        _ck(self.lib, self.lib.vb_download_factor_cache(self.h, store, _d(out), out.size))
        return out

    def store_aux(self, store):
        out = np.zeros(self.g.stores[store].aux.size)
        _ck(self.lib, self.lib.vb_download_store_aux(self.h, store, _d(out), out.size))
        return out

    def rs_set_imu(self, store, time_ns, gyro, accel, rig_mid_time_us, rig_calib_var, half_length_us):
        """raw IMU-0 samples + per-rig windows of a rolling-shutter store (vb_rs_set_imu): from now on optimize()
        recomputes the store's tables from the current calibration at the top of every iteration"""
        t = np.ascontiguousarray(time_ns, np.int64); gy = np.ascontiguousarray(gyro, np.float32); ac = np.ascontiguousarray(accel, np.float32)
        mid = np.ascontiguousarray(rig_mid_time_us, np.int64); cv = np.ascontiguousarray(rig_calib_var, np.int32)
        assert gy.shape == (t.size, 3) and ac.shape == (t.size, 3) and mid.size == cv.size
        _ck(self.lib, self.lib.vb_rs_set_imu(self.h, int(store), mid.size, mid.ctypes.data_as(_ip), cv.ctypes.data_as(C.POINTER(C.c_int32)),
                                             int(half_length_us), t.size, t.ctypes.data_as(_ip), gy.ctypes.data_as(C.POINTER(C.c_float)),
                                             ac.ctypes.data_as(C.POINTER(C.c_float))))

    def rs_clear_imu(self, store):
        _ck(self.lib, self.lib.vb_rs_set_imu(self.h, int(store), 0, None, None, 0, 0, None, None, None))

    def rs_refresh(self, store=-1):
        _ck(self.lib, self.lib.vb_rs_refresh(self.h, int(store)))

    def set_linear_solver(self, solver_type, max_iterations=40, desired_residual=1e-10):
        """Settings::solverType for vb_factor / vb_solve (optimize() takes it from its settings): graph.SOLVER_DIRECT,
        SOLVER_PCG_TRIVIAL, SOLVER_PCG_JACOBI, SOLVER_PCG_GAUSS_SEIDEL, SOLVER_PCG_LOWER_PREC_SOLVE"""
        _ck(self.lib, self.lib.vb_set_linear_solver(self.h, int(solver_type), int(max_iterations), float(desired_residual)))

    def pcg_last(self):
        """(iterations, relative residual) of the last PCG solve (PCG::solve's return, PCG.cpp:95)"""
        it = C.c_int32(0); res = C.c_double(0.0)
        _ck(self.lib, self.lib.vb_pcg_last(self.h, C.byref(it), C.byref(res)))
        return it.value, res.value

    def rs_set_gravity_var(self, store, gravity_var):
        _ck(self.lib, self.lib.vb_rs_set_gravity_var(self.h, int(store), int(gravity_var)))

    def preint_set_imu(self, store, time_ns, gyro, accel, start_us, end_us, calib_var, gyro_var, accel_var):
        """raw samples of an inertial store's IMU + the interval and calibration variable of every factor (vb_preint_set_imu):
        from now on optimize() recomputes the store's preintegration records at the top of every iteration"""
        t = np.ascontiguousarray(time_ns, np.int64); gy = np.ascontiguousarray(gyro, np.float32); ac = np.ascontiguousarray(accel, np.float32)
        t0 = np.ascontiguousarray(start_us, np.int64); t1 = np.ascontiguousarray(end_us, np.int64); cv = np.ascontiguousarray(calib_var, np.int32)
        gv = np.ascontiguousarray(gyro_var, np.float64); av = np.ascontiguousarray(accel_var, np.float64)
        assert gy.shape == (t.size, 3) and ac.shape == (t.size, 3) and t0.size == t1.size == cv.size and gv.size == 3 and av.size == 3
        fp = C.POINTER(C.c_float)
        self.lib.vb_preint_set_imu.argtypes = [C.c_void_p, C.c_int32, C.c_int64, _ip, _ip, C.POINTER(C.c_int32), C.c_int64, _ip, fp, fp, _dp, _dp]
        _ck(self.lib, self.lib.vb_preint_set_imu(self.h, int(store), cv.size, t0.ctypes.data_as(_ip), t1.ctypes.data_as(_ip),
                                                 cv.ctypes.data_as(C.POINTER(C.c_int32)), t.size, t.ctypes.data_as(_ip), gy.ctypes.data_as(fp),
                                                 ac.ctypes.data_as(fp), _d(gv), _d(av)))

    def preint_clear_imu(self, store):
        self.lib.vb_preint_set_imu.argtypes = [C.c_void_p, C.c_int32, C.c_int64, _ip, _ip, C.POINTER(C.c_int32), C.c_int64, _ip,
                                               C.POINTER(C.c_float), C.POINTER(C.c_float), _dp, _dp]
        _ck(self.lib, self.lib.vb_preint_set_imu(self.h, int(store), 0, None, None, None, 0, None, None, None, None, None))

    def preint_refresh(self, store=-1):
        _ck(self.lib, self.lib.vb_preint_refresh(self.h, int(store)))

    def preint_download(self, store, n_factors):
        cov, om = np.zeros((n_factors, 9, 9)), np.zeros((n_factors, 3))
        self.lib.vb_preint_download.argtypes = [C.c_void_p, C.c_int32, _dp, _dp, C.c_int64]
        _ck(self.lib, self.lib.vb_preint_download(self.h, int(store), _d(cov), _d(om), n_factors))
        return cov, om

    def joint_covariance(self, indices, damping=1e-5):
        """Optimizer::computeJointCovariances for one list of registration indices: (cov [D, D], offsets [n + 1], damping used)"""
        idx = np.ascontiguousarray(indices, np.int64)
        dims = [int(self.g.var_tdim[np.flatnonzero(self.g.var_index == i)[0]]) for i in idx]
        D = int(sum(dims))
        cov, off, used = np.zeros((D, D)), np.zeros(idx.size + 1, np.int64), C.c_double(0)
        self.lib.vb_compute_joint_covariance.argtypes = [C.c_void_p, C.c_double, C.c_int64, _ip, _dp, C.c_int64, _ip, C.POINTER(C.c_double)]
        _ck(self.lib, self.lib.vb_compute_joint_covariance(self.h, damping, idx.size, idx.ctypes.data_as(_ip), _d(cov), D * D,
                                                           off.ctypes.data_as(_ip), C.byref(used)))
        return cov, off, used.value

    def joint_covariances(self, lists, damping=1e-5):
        """Optimizer::computeJointCovariances for several lists of registration indices, one factorisation: ([cov_l], damping used)"""
        ptr = np.concatenate([[0], np.cumsum([len(l) for l in lists])]).astype(np.int64)
        idx = np.ascontiguousarray(np.concatenate([np.asarray(l, np.int64) for l in lists]), np.int64)
        tdim = {int(i): int(t) for i, t in zip(self.g.var_index, self.g.var_tdim) if i >= 0}
        D = [sum(tdim[int(i)] for i in l) for l in lists]
        total = int(sum(d * d for d in D))
        cov, cptr, used = np.zeros(total), np.zeros(len(lists) + 1, np.int64), C.c_double(0)
        self.lib.vb_compute_joint_covariances.argtypes = [C.c_void_p, C.c_double, C.c_int64, _ip, _ip, _dp, C.c_int64, _ip, C.POINTER(C.c_double)]
        _ck(self.lib, self.lib.vb_compute_joint_covariances(self.h, damping, len(lists), ptr.ctypes.data_as(_ip), idx.ctypes.data_as(_ip), _d(cov),
                                                            total, cptr.ctypes.data_as(_ip), C.byref(used)))
        return [cov[cptr[l]:cptr[l + 1]].reshape(D[l], D[l]).copy() for l in range(len(lists))], used.value

    def marginal_problem(self, indices, damping=1e-5):
        """Optimizer::computeMarginalProblem for one list of registration indices: (H [D, D], b [D], cost, offsets, damping used)"""
        idx = np.ascontiguousarray(indices, np.int64)
        dims = [int(self.g.var_tdim[np.flatnonzero(self.g.var_index == i)[0]]) for i in idx]
        D = int(sum(dims))
        H, b, off = np.zeros((D, D)), np.zeros(D), np.zeros(idx.size + 1, np.int64)
        cost, used = C.c_double(0), C.c_double(0)
        self.lib.vb_compute_marginal_problem.argtypes = [C.c_void_p, C.c_double, C.c_int64, _ip, _dp, C.c_int64, _dp, C.c_int64,
                                                         C.POINTER(C.c_double), _ip, C.POINTER(C.c_double)]
        _ck(self.lib, self.lib.vb_compute_marginal_problem(self.h, damping, idx.size, idx.ctypes.data_as(_ip), _d(H), D * D, _d(b), D,
                                                           C.byref(cost), off.ctypes.data_as(_ip), C.byref(used)))
        return H, b, cost.value, off, used.value

    def refine_points(self, per_point=True):
        """refinePoints pre-pass (vb_refine_points): returns (stats dict, per-point array [n, 3] = startCost, endCost, nIts)"""
        class _S(C.Structure):
            _fields_ = [("total_start_cost", C.c_double), ("total_end_cost", C.c_double), ("num_points", C.c_int64),
                        ("num_failures", C.c_int64), ("num_total_iterations", C.c_int64), ("num_at_least_one_iteration", C.c_int64)]
        self.lib.vb_refine_num_points.restype = C.c_int64
        self.lib.vb_refine_num_points.argtypes = [C.c_void_p]
        n = self.lib.vb_refine_num_points(self.h)
        pp = np.zeros((n, 3)) if per_point else None
        st = _S()
        self.lib.vb_refine_points.argtypes = [C.c_void_p, C.POINTER(_S), _dp, C.c_int64]
        _ck(self.lib, self.lib.vb_refine_points(self.h, C.byref(st), _d(pp) if per_point else None, n))
        return {k: getattr(st, k) for k, _ in _S._fields_}, pp

    def refine_triangulation(self, outlier_threshold, skip_outliers, max_num_iterations, huber_radius):
        """refineTriangulationResult over every landmark (vb_refine_triangulation): (per-point [n, 3] = numInliers, last
        cost, status; inlier flags of the visual stores concatenated in graph order; failures)"""
        self.lib.vb_refine_num_points.restype = C.c_int64
        self.lib.vb_refine_num_points.argtypes = [C.c_void_p]
        n = self.lib.vb_refine_num_points(self.h)
        pp = np.zeros((n, 3)); fails = C.c_int64(0)
        self.lib.vb_refine_triangulation.argtypes = [C.c_void_p, C.c_double, C.c_int32, C.c_int32, C.c_double, _dp, C.c_int64, C.POINTER(C.c_int64)]
        _ck(self.lib, self.lib.vb_refine_triangulation(self.h, float(outlier_threshold), int(bool(skip_outliers)), int(max_num_iterations),
                                                       float(huber_radius), _d(pp), n, C.byref(fails)))
        self.lib.vb_refine_triangulation_inliers.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_uint8), C.c_int64]
        flags = []
        for si, st in enumerate(self.g.stores):
            if st.type not in (G.F_VISUAL, G.F_VISUAL_RS):
                continue
            f = np.zeros(st.vars.shape[0], np.uint8)
            _ck(self.lib, self.lib.vb_refine_triangulation_inliers(self.h, si, f.ctypes.data_as(C.POINTER(C.c_uint8)), f.size))
            flags.append(f)
        return pp, (np.concatenate(flags) if flags else np.zeros(0, np.uint8)), fails.value

    def upload_store_aux(self, store, aux):
        aux = np.ascontiguousarray(aux, np.float64)
        _ck(self.lib, self.lib.vb_upload_store_aux(self.h, store, _d(aux), aux.size))
