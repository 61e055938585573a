"""Host-side graph description: numpy SoA arrays <-> the C ABI's ``vb_graph_desc``.

Mirrors what the reference's ``small_thing::Optimizer`` holds after all ``addFactor`` calls
(lib/small_thing/Optimizer.h:115-175, Factor.h:96-141): variables by kind, one table per
factor store, and the registration rule of ``registerAllVariables``
(lib/small_thing/Optimizer.cpp:25-39, Factor.h:737-758).
"""
import ctypes as C
import numpy as np

# VB_VAR_*
VAR_VEC3, VAR_SE3, VAR_S2, VAR_CAM, VAR_IMUCALIB, VAR_VEC2 = range(6)
DATA_DIM = {VAR_VEC3: 3, VAR_SE3: 7, VAR_S2: 3, VAR_CAM: 17, VAR_IMUCALIB: 32, VAR_VEC2: 2}
CAM_EST_READOUT, CAM_EST_TIMEOFFSET, CAM_HAS_READOUT = 1, 2, 4
# VB_F_*
F_VISUAL, F_VISUAL_RS = 1, 2
F_INERTIAL, F_INERTIAL_SEC, F_INERTIAL_SEC_SPLIT = 4, 5, 6
F_RW_IMUCALIB, F_RW_CAM, F_RW_SE3 = 7, 8, 9
F_PRIOR_IMUCALIB, F_PRIOR_CAM, F_PRIOR_SE3 = 11, 12, 13
F_OMEGA_PRIOR, F_OMEGA_PRIOR_SEC = 15, 16
ERR_SIZE = {1: 2, 2: 2, 4: 9, 5: 9, 6: 9, 7: 23, 8: 17, 9: 6, 11: 23, 12: 17, 13: 6, 15: 3, 16: 3}
PREINT_STRIDE = 250  # R(4) dV(3) dP(3) dt(1) J(9x23) evalPoint(32)


class vb_factor_store(C.Structure):
    _fields_ = [("type", C.c_int32), ("nvars", C.c_int32), ("count", C.c_int64),
                ("vars", C.POINTER(C.c_int32)), ("consts_per_factor", C.c_int32),
                ("consts", C.POINTER(C.c_double)), ("loss_a", C.c_double), ("loss_k", C.c_double),
                ("aux_len", C.c_int64), ("aux", C.POINTER(C.c_double))]


class vb_graph_desc(C.Structure):
    _fields_ = [("num_vars", C.c_int64), ("var_kind", C.POINTER(C.c_int32)),
                ("var_tdim", C.POINTER(C.c_int32)), ("var_flags", C.POINTER(C.c_uint32)),
                ("var_index", C.POINTER(C.c_int64)), ("var_off", C.POINTER(C.c_int64)),
                ("var_aux", C.POINTER(C.c_double)), ("var_data_len", C.c_int64),
                ("var_data", C.POINTER(C.c_double)), ("num_points", C.c_int64),
                ("num_stores", C.c_int32), ("stores", C.POINTER(vb_factor_store))]


class vb_settings(C.Structure):
    _fields_ = [(n, C.c_double) for n in (
        "damping", "damping_adjust_on_fail", "damping_adjust_on_good", "damping_adjust_on_average",
        "damping_min", "damping_max", "min_relative_cost_reduction", "step_factor_decrease",
        "min_step_factor_for_good", "absolute_cost_tolerance", "relative_cost_tolerance",
        "variables_tolerance")] + [(n, C.c_int32) for n in (
        "max_step_factor_attempts", "try_sub_step", "stop_if_no_improvement_for",
        "distance_from_troubled_iteration", "max_num_iterations", "verbose",
        "solver_type", "pcg_max_iterations")] + [("pcg_desired_residual", C.c_double)]


class vb_summary(C.Structure):
    _fields_ = [("initial_cost", C.c_double), ("final_cost", C.c_double),
                ("num_troubled_seqs", C.c_int32), ("largest_troubled_seq", C.c_int32),
                ("num_iterations", C.c_int32), ("stop_reason", C.c_int32)]


class vb_cost_stats(C.Structure):
    _fields_ = [("num_total", C.c_int64), ("num_invalid", C.c_int64), ("num_prev_invalid", C.c_int64)]


class vb_iter_times(C.Structure):
    _fields_ = [(n, C.c_double) for n in ("grad_hess", "damping", "factor_schur", "factor_chol",
                                          "solve", "apply", "cost", "total", "step_dot_grad")]


# VB_BR_* bits of vb_iter_trace.branch (include/viba_b200.h)
BR_MODEL_NEGATIVE, BR_STEP_FACTOR, BR_ACCEPT_AT_FACTOR, BR_SUB_STEP, BR_ACCEPT_AT_SUB = 1, 2, 4, 8, 16
BR_LATCH_NO_RETRY, BR_REJECTED, BR_GOOD, BR_AVERAGE, BR_TOLERANCE_HIT, BR_TROUBLED_SEQ_END = 32, 64, 128, 256, 512, 1024


class vb_iter_trace(C.Structure):
    _fields_ = [(n, C.c_double) for n in (
        "prev_cost", "new_cost", "damping", "damping_used", "model_cost_reduction", "applied_step_factor",
        "ratio_reduction_to_expected", "cost_reduction", "grad_norm", "step_norm",
        "ratio_s2v_inf", "ratio_s2v_2", "ratio_s2v_1", "ms_total", "ms_grad_hess", "ms_factor", "ms_solve")] + [
        (n, C.c_int64) for n in ("num_total", "num_invalid", "num_prev_invalid")] + [
        ("branch", C.c_int32), ("step_factor_attempts", C.c_int32)]


PRE_STEP_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int32)
LOG_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_char_p)


class vb_optimize_hooks(C.Structure):
    _fields_ = [("pre_step", PRE_STEP_FN), ("pre_step_user", C.c_void_p), ("sync_kinds", C.c_uint32),
                ("host_vars", C.POINTER(C.c_double)), ("log", LOG_FN), ("log_user", C.c_void_p),
                ("trace", C.POINTER(vb_iter_trace)), ("trace_cap", C.c_int32), ("trace_len", C.c_int32)]


def trace_records(arr, n):
    """list of dicts from the first n vb_iter_trace records"""
    return [{name: getattr(arr[i], name) for name, _ in vb_iter_trace._fields_} for i in range(n)]


SOLVER_DIRECT, SOLVER_PCG_TRIVIAL, SOLVER_PCG_JACOBI, SOLVER_PCG_GAUSS_SEIDEL, SOLVER_PCG_LOWER_PREC_SOLVE = range(5)


def default_settings(max_iterations=250):
    """Optimizer::Settings defaults (lib/small_thing/Optimizer.h:40-78) with the ark_vi_ba
    overrides (viba/common/Settings.cpp:313-320); SURVEY.md App. A."""
    return vb_settings(1e-5, 2.5, 0.7, 1.5, 1e-9, 1e8, 0.3, 0.3, 0.7, 1e-8, 1e-10, 1e-5,
                       2, 1, 3, 3, max_iterations, 0, SOLVER_DIRECT, 40, 1e-10)


class FactorStore:
    def __init__(self, ftype, vars_, consts, loss=(0.0, 0.0), aux=None):
        self.type = int(ftype)
        self.vars = np.ascontiguousarray(vars_, dtype=np.int32)
        assert self.vars.ndim == 2
        self.consts = np.ascontiguousarray(consts, dtype=np.float64).reshape(self.vars.shape[0], -1)
        self.loss = (float(loss[0]), float(loss[1]))
        self.aux = None if aux is None else np.ascontiguousarray(aux, dtype=np.float64).ravel()

    @property
    def count(self):
        return self.vars.shape[0]


class Graph:
    """Variable pool + factor stores.  ``add_vars`` returns pool ids; ``finalize`` assigns the
    registration indices the way the reference does."""

    def __init__(self):
        self._kind, self._tdim, self._flags, self._aux, self._data, self._const = [], [], [], [], [], []
        self.stores = []
        self.num_points = 0
        self.nv = 0

    def add_vars(self, kind, data, tdim, flags=0, aux=0.0, constant=False):
        data = np.asarray(data, dtype=np.float64).reshape(-1, DATA_DIM[kind])
        n = data.shape[0]
        self._kind.append(np.full(n, kind, np.int32))
        self._tdim.append(np.broadcast_to(np.asarray(tdim, np.int32), (n,)).copy())
        self._flags.append(np.broadcast_to(np.asarray(flags, np.uint32), (n,)).copy())
        self._aux.append(np.broadcast_to(np.asarray(aux, np.float64), (n,)).copy())
        self._const.append(np.full(n, bool(constant)))
        self._data.append(data.ravel())
        ids = np.arange(self.nv, self.nv + n, dtype=np.int32)
        self.nv += n
        return ids

    def add_store(self, store):
        self.stores.append(store)
        return len(self.stores) - 1

    def finalize(self, point_ids):
        """registerPointVariables + registeredVariablesToEliminationRange
        (viba/problem/SingleSessionProblem.cpp:41-45, main_AriaKit_ViBa.cpp:76-77) then
        registerAllVariables: per store, per argument position, per factor in insertion order."""
        self.var_kind = np.concatenate(self._kind)
        self.var_tdim = np.concatenate(self._tdim)
        self.var_flags = np.concatenate(self._flags)
        self.var_aux = np.concatenate(self._aux)
        self.var_data = np.concatenate(self._data)
        const = np.concatenate(self._const)
        dd = np.array([DATA_DIM[k] for k in range(6)], np.int64)[self.var_kind]
        self.var_off = np.concatenate([[0], np.cumsum(dd)[:-1]]).astype(np.int64)
        idx = np.full(self.nv, -1, np.int64)
        idx[const] = -2
        point_ids = np.asarray(point_ids, np.int64)
        idx[point_ids] = np.arange(len(point_ids))
        nxt = len(point_ids)
        self.num_points = len(point_ids)
        for st in self.stores:
            for a in range(st.vars.shape[1]):
                col = st.vars[:, a]
                un = col[idx[col] == -1]
                if un.size == 0:
                    continue
                _, first = np.unique(un, return_index=True)
                new = un[np.sort(first)]
                idx[new] = nxt + np.arange(new.size)
                nxt += new.size
        self.var_index = idx
        self.num_blocks = nxt
        return self

    # ---- ctypes view ---------------------------------------------------------------------
    def desc(self):
        def p(a, t):
            return a.ctypes.data_as(C.POINTER(t))
        self._keep = []
        arr = (vb_factor_store * len(self.stores))()
        for i, st in enumerate(self.stores):
            a = arr[i]
            a.type, a.nvars, a.count = st.type, st.vars.shape[1], st.count
            a.vars = p(st.vars, C.c_int32)
            a.consts_per_factor = st.consts.shape[1]
            a.consts = p(st.consts, C.c_double)
            a.loss_a, a.loss_k = st.loss
            if st.aux is not None:
                a.aux_len, a.aux = st.aux.size, p(st.aux, C.c_double)
            else:
                a.aux_len, a.aux = 0, None
        self._keep.append(arr)
        d = vb_graph_desc()
        d.num_vars = self.nv
        d.var_kind = p(self.var_kind, C.c_int32)
        d.var_tdim = p(self.var_tdim, C.c_int32)
        d.var_flags = p(self.var_flags, C.c_uint32)
        d.var_index = p(self.var_index, C.c_int64)
        d.var_off = p(self.var_off, C.c_int64)
        d.var_aux = p(self.var_aux, C.c_double)
        d.var_data_len = self.var_data.size
        d.var_data = p(self.var_data, C.c_double)
        d.num_points = self.num_points
        d.num_stores = len(self.stores)
        d.stores = arr
        return d

    def num_factors(self):
        return sum(s.count for s in self.stores)
