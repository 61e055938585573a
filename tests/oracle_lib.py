"""ctypes binding of oracle/liboracle.so (the CPU checker).  Test infrastructure only."""
import ctypes as C
import os, subprocess
import numpy as np
from visual_inertial_bundle_adjustment_b200 import graph as G

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int64)


def load_oracle():
    so = os.path.join(ROOT, "oracle", "liboracle.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
    lib = C.CDLL(so)
    lib.orc_last_error.restype = C.c_char_p
    for f in ("orc_order", "orc_data_size", "orc_num_blocks", "orc_nnz_blocks"):
        getattr(lib, f).restype = C.c_int64
        getattr(lib, f).argtypes = [C.c_void_p]
    lib.orc_damp.argtypes = [C.c_void_p, C.c_double]
    lib.orc_scale_vec.argtypes = [C.c_void_p, C.c_int, C.c_double]
    lib.orc_boxplus.argtypes = [C.c_int, C.c_uint32, C.c_double, _dp, _dp, C.c_int, _dp]
    lib.orc_lm_iteration_timed.argtypes = [C.c_void_p, C.c_double, _dp, _dp, _dp]
    lib.orc_eval_factor.argtypes = [C.c_void_p, C.c_int, C.c_int64, _dp, _dp, C.POINTER(C.c_int)]
    lib.orc_set_parallel.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_char_p]
    return lib


def find_openblas():
    """scipy's bundled OpenBLAS (LAPACK symbols prefixed `scipy_`), or None."""
    import glob
    try:
        import scipy
        hits = glob.glob(os.path.join(os.path.dirname(os.path.dirname(scipy.__file__)), "scipy.libs", "libscipy_openblas*.so"))
        return (hits[0], "scipy_") if hits else None
    except Exception:
        return None


def dptr(a):
    return a.ctypes.data_as(_dp)


class Oracle:
    def __init__(self, lib, graph):
        self.lib, self.g = lib, graph
        self.h = C.c_void_p()
        self._desc = graph.desc()
        self._ck(lib.orc_create(C.byref(self._desc), C.byref(self.h)))
        self.order = lib.orc_order(self.h)
        self.data_size = lib.orc_data_size(self.h)
        self.nblocks = lib.orc_num_blocks(self.h)
        self.nnz = lib.orc_nnz_blocks(self.h)

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError(self.lib.orc_last_error().decode())

    def __del__(self):
        try:
            self.lib.orc_destroy(self.h)
        except Exception:
            pass

    def set_parallel(self, threads):
        """CPU-baseline mode: OpenMP factor loops + landmark elimination on `threads` threads and, when
        scipy's OpenBLAS is present, LAPACK dpbtrf/dpbtrs for the reduced system.  Returns the solver kind."""
        blas = find_openblas() if threads > 1 else None
        self._ck(self.lib.orc_set_parallel(self.h, int(threads), blas[0].encode() if blas else None,
                                           blas[1].encode() if blas else None))
        return "openblas-dpbtrf" if blas else "envelope"

    def structure(self):
        rp = np.zeros(self.nblocks + 1, np.int64); ci = np.zeros(self.nnz, np.int64)
        bo = np.zeros(self.nnz, np.int64); ps = np.zeros(self.nblocks + 1, np.int64)
        self.lib.orc_get_structure(self.h, rp.ctypes.data_as(_ip), ci.ctypes.data_as(_ip), bo.ctypes.data_as(_ip), ps.ctypes.data_as(_ip))
        return rp, ci, bo, ps

    def eval_grad_hess(self, with_hess=True, update_cache=True, dont_retry=False, which=0):
        c = C.c_double()
        self._ck(self.lib.orc_eval_grad_hess(self.h, int(with_hess), int(update_cache), int(dont_retry), which, C.byref(c)))
        return c.value

    def eval_cost(self, comparable=True):
        c = C.c_double(); st = G.vb_cost_stats()
        self._ck(self.lib.orc_eval_cost(self.h, int(comparable), C.byref(c), C.byref(st)))
        return c.value, (st.num_total, st.num_invalid, st.num_prev_invalid)

    def damp(self, lam): self.lib.orc_damp(self.h, lam)
    def factor(self): self._ck(self.lib.orc_factor(self.h))
    def solve(self, which=0): self._ck(self.lib.orc_solve(self.h, which))
    def scale(self, which, s): self.lib.orc_scale_vec(self.h, which, s)
    def backup(self): self.lib.orc_backup(self.h)
    def restore(self): self.lib.orc_restore(self.h)

    def apply_step(self, which=0):
        r = np.zeros(3)
        self.lib.orc_apply_step(self.h, which, dptr(r))
        return r

    def get(self, which):
        n = {0: self.order, 1: self.order, 2: self.order, 3: self.data_size, 4: self.g.var_data.size}[which]
        out = np.zeros(n)
        self._ck(self.lib.orc_get(self.h, which, dptr(out), n))
        return out

    def set(self, which, a):
        a = np.ascontiguousarray(a, np.float64)
        self._ck(self.lib.orc_set(self.h, which, dptr(a), a.size))

    grad = property(lambda s: s.get(0)); step = property(lambda s: s.get(1))
    hess = property(lambda s: s.get(3)); vars = property(lambda s: s.get(4))

    def rs_set_imu(self, store, time_ns, gyro, accel, rig_mid_time_us, rig_calib_var, half_length_us):
        t = np.ascontiguousarray(time_ns, np.int64); gy = np.ascontiguousarray(gyro, np.float32); ac = np.ascontiguousarray(accel, np.float32)
        mid = np.ascontiguousarray(rig_mid_time_us, np.int64); cv = np.ascontiguousarray(rig_calib_var, np.int32)
        fp = C.POINTER(C.c_float)
        self.lib.orc_rs_set_imu.argtypes = [C.c_void_p, C.c_int, C.c_int64, _ip, C.POINTER(C.c_int32), C.c_int64, C.c_int64, _ip, fp, fp]
        self._ck(self.lib.orc_rs_set_imu(self.h, int(store), mid.size, mid.ctypes.data_as(_ip), cv.ctypes.data_as(C.POINTER(C.c_int32)),
                                         int(half_length_us), t.size, t.ctypes.data_as(_ip), gy.ctypes.data_as(fp), ac.ctypes.data_as(fp)))

    def preint_set_imu(self, store, time_ns, gyro, accel, start_us, end_us, calib_var, gyro_var, accel_var):
        t = np.ascontiguousarray(time_ns, np.int64); gy = np.ascontiguousarray(gyro, np.float32); ac = np.ascontiguousarray(accel, np.float32)
        t0 = np.ascontiguousarray(start_us, np.int64); t1 = np.ascontiguousarray(end_us, np.int64); cv = np.ascontiguousarray(calib_var, np.int32)
        gv = np.ascontiguousarray(gyro_var, np.float64); av = np.ascontiguousarray(accel_var, np.float64)
        fp = C.POINTER(C.c_float)
        self.lib.orc_preint_set_imu.argtypes = [C.c_void_p, C.c_int, C.c_int64, _ip, _ip, C.POINTER(C.c_int32), C.c_int64, _ip, fp, fp, _dp, _dp]
        self._ck(self.lib.orc_preint_set_imu(self.h, int(store), cv.size, t0.ctypes.data_as(_ip), t1.ctypes.data_as(_ip), cv.ctypes.data_as(C.POINTER(C.c_int32)),
                                             t.size, t.ctypes.data_as(_ip), gy.ctypes.data_as(fp), ac.ctypes.data_as(fp), dptr(gv), dptr(av)))

    def preint_refresh(self, store=-1):
        self.lib.orc_preint_refresh.argtypes = [C.c_void_p, C.c_int]
        self._ck(self.lib.orc_preint_refresh(self.h, int(store)))

    def preint_download(self, store, n_factors):
        cov, om = np.zeros((n_factors, 9, 9)), np.zeros((n_factors, 3))
        self.lib.orc_preint_get_cov.argtypes = [C.c_void_p, C.c_int, _dp, _dp, C.c_int64]
        self._ck(self.lib.orc_preint_get_cov(self.h, int(store), dptr(cov), dptr(om), n_factors))
        return cov, om

    def set_linear_solver(self, solver_type, max_iterations=40, residual=1e-10):
        self.lib.orc_set_linear_solver.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double]
        self._ck(self.lib.orc_set_linear_solver(self.h, int(solver_type), int(max_iterations), float(residual)))

    def pcg_last(self):
        it, res = C.c_int(0), C.c_double(0)
        self.lib.orc_pcg_last.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_double)]
        self.lib.orc_pcg_last(self.h, C.byref(it), C.byref(res))
        return it.value, res.value

    @staticmethod
    def pcg_dense(lib, A, b, block_size, precond, desired_residual, max_iterations):
        """PCG::solve on a dense symmetric system with IdentityPrecond (0) / BlockJacobiPrecond (1): (x, iterations, residual)"""
        A = np.ascontiguousarray(A, np.float64); b = np.ascontiguousarray(b, np.float64)
        bs = np.ascontiguousarray(block_size, np.int32)
        x = np.zeros(b.size); it = C.c_int(0); res = C.c_double(0)
        lib.orc_pcg_dense.argtypes = [C.c_int64, _dp, _dp, C.POINTER(C.c_int32), C.c_int64, C.c_int, C.c_double, C.c_int, _dp,
                                      C.POINTER(C.c_int), C.POINTER(C.c_double)]
        rc = lib.orc_pcg_dense(b.size, dptr(A), dptr(b), bs.ctypes.data_as(C.POINTER(C.c_int32)), bs.size, int(precond),
                               float(desired_residual), int(max_iterations), dptr(x), C.byref(it), C.byref(res))
        if rc != 0:
            raise RuntimeError("orc_pcg_dense failed")
        return x, it.value, res.value

    def refine_points(self):
        """refinePoints on the oracle: (stats dict, per-point [n, 3]); the landmark variables are updated"""
        n = len(set(int(v) for st in self.g.stores if st.type in (G.F_VISUAL, G.F_VISUAL_RS) for v in st.vars[:, 0]))
        pp, st = np.zeros((n, 3)), np.zeros(6)
        self.lib.orc_refine_points.argtypes = [C.c_void_p, _dp, _dp]
        self._ck(self.lib.orc_refine_points(self.h, dptr(st), dptr(pp)))
        keys = ["total_start_cost", "total_end_cost", "num_points", "num_failures", "num_total_iterations", "num_at_least_one_iteration"]
        return {k: (float(v) if i < 2 else int(v)) for i, (k, v) in enumerate(zip(keys, st))}, pp

    def refine_triangulation(self, outlier_threshold, skip_outliers, max_num_iterations, huber_radius):
        """refineTriangulationResult over every landmark: (per-point [n, 3] = numInliers, last cost, status; inlier flags of
        the visual stores concatenated in graph order; failures); the landmark variables are updated"""
        vis = [st for st in self.g.stores if st.type in (G.F_VISUAL, G.F_VISUAL_RS)]
        n = len(set(int(v) for st in vis for v in st.vars[:, 0]))
        pp = np.zeros((n, 3)); fl = np.zeros(sum(st.vars.shape[0] for st in vis), np.uint8); fails = C.c_int64(0)
        self.lib.orc_refine_triangulation.argtypes = [C.c_void_p, C.c_double, C.c_int, C.c_int, C.c_double, _dp, C.POINTER(C.c_uint8),
                                                      C.POINTER(C.c_int64)]
        self._ck(self.lib.orc_refine_triangulation(self.h, float(outlier_threshold), int(bool(skip_outliers)), int(max_num_iterations),
                                                   float(huber_radius), dptr(pp), fl.ctypes.data_as(C.POINTER(C.c_uint8)), C.byref(fails)))
        return pp, fl, fails.value

    def rs_set_gravity_var(self, store, gravity_var):
        self.lib.orc_rs_set_gravity_var.argtypes = [C.c_void_p, C.c_int, C.c_int]
        self._ck(self.lib.orc_rs_set_gravity_var(self.h, int(store), int(gravity_var)))

    def rs_refresh(self, store=-1):
        self.lib.orc_rs_refresh.argtypes = [C.c_void_p, C.c_int]
        self._ck(self.lib.orc_rs_refresh(self.h, int(store)))

    def store_aux(self, store):
        out = np.zeros(self.g.stores[store].aux.size)
        self.lib.orc_get_store_aux.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.c_int64]
        assert self.lib.orc_get_store_aux(self.h, int(store), dptr(out), out.size) == 0
        return out

    def cache(self, store):
        out = np.zeros(self.g.stores[store].count)
        self.lib.orc_get_cache(self.h, store, dptr(out), out.size)
        return out

    def eval_factor(self, store, k, jac=True):
        st = self.g.stores[store]
        E = G.ERR_SIZE[st.type]
        dims = self.g.var_tdim[st.vars[k]]
        r = np.zeros(E); J = np.zeros(E * int(dims.sum())); valid = C.c_int()
        self._ck(self.lib.orc_eval_factor(self.h, store, k, dptr(r), dptr(J) if jac else None, C.byref(valid)))
        Js, o = [], 0
        for d in dims:
            Js.append(J[o:o + E * d].reshape(E, d)); o += E * d
        return r, Js, bool(valid.value)

    def optimize(self, settings, trace_cap=300):
        summ = G.vb_summary(); tp = np.zeros(trace_cap); tn = np.zeros(trace_cap)
        self._ck(self.lib.orc_optimize(self.h, C.byref(settings), C.byref(summ), dptr(tp), dptr(tn), trace_cap))
        n = summ.num_iterations
        return summ, tp[:n], tn[:n]

    def optimize_ex(self, settings, trace_cap=300, pre_step=None):
        """optimize with the full per-iteration trace (vb_iter_trace records) and an optional
        pre_step(iteration) callback (Settings::preStepCallback stand-in)"""
        summ = G.vb_summary(); tr = (G.vb_iter_trace * trace_cap)(); n = C.c_int()
        CB = C.CFUNCTYPE(None, C.c_void_p, C.c_int32)
        cb = CB((lambda user, it: pre_step(it)) if pre_step else 0)
        self.lib.orc_optimize_ex.argtypes = [C.c_void_p, C.POINTER(G.vb_settings), C.POINTER(G.vb_summary),
                                             C.POINTER(G.vb_iter_trace), C.c_int, C.POINTER(C.c_int), CB, C.c_void_p]
        self._ck(self.lib.orc_optimize_ex(self.h, C.byref(settings), C.byref(summ), tr, trace_cap, C.byref(n), cb, None))
        return summ, G.trace_records(tr, n.value)

    def set_store_aux(self, store, aux):
        aux = np.ascontiguousarray(aux, np.float64)
        self.lib.orc_set_store_aux.argtypes = [C.c_void_p, C.c_int, _dp, C.c_int64]
        self._ck(self.lib.orc_set_store_aux(self.h, int(store), dptr(aux), aux.size))

    def lm_iteration_timed(self, lam):
        """ms[0..6] = g/H, damping, factor, solve, apply, cost, total; ms[7] = step.dot(grad)"""
        ms = np.zeros(8); c0 = C.c_double(); c1 = C.c_double()
        self._ck(self.lib.orc_lm_iteration_timed(self.h, lam, dptr(ms), C.byref(c0), C.byref(c1)))
        return ms, c0.value, c1.value

    def dense_hessian(self):
        """Symmetric dense H in registration order from the block CSR (small problems only)."""
        rp, ci, bo, ps = self.structure()
        H = np.zeros((self.order, self.order)); h = self.hess
        for i in range(self.nblocks):
            di = ps[i + 1] - ps[i]
            for k in range(rp[i], rp[i + 1]):
                j = ci[k]; dj = ps[j + 1] - ps[j]
                B = h[bo[k]:bo[k] + di * dj].reshape(di, dj)
                H[ps[i]:ps[i + 1], ps[j]:ps[j + 1]] = B
                if i != j:
                    H[ps[j]:ps[j + 1], ps[i]:ps[i + 1]] = B.T
        return H
