"""Oracle self-verification: every factor's analytic Jacobian against central differences through
applyBoxPlus (the recipe of Optimizer::verifyJacobians, lib/small_thing/Factor.h:256-387, made
tighter), the assembled gradient against finite differences of the cost, factor+solve against a
dense solve, and LM convergence — the identities the reference's own tests use."""
import ctypes as C
import numpy as np
import pytest
from tests.oracle_lib import Oracle, dptr
from visual_inertial_bundle_adjustment_b200 import synth, graph as G


@pytest.fixture(scope="module", params=["tiny", "tiny_rs"])
def tiny(request, oracle):
    s = synth.make_session(**synth.CONFIGS[request.param])
    return s, Oracle(oracle, s.graph)


def _boxplus(lib, g, v, x, s):
    x = x.copy()
    lib.orc_boxplus(int(g.var_kind[v]), int(g.var_flags[v]), float(g.var_aux[v]), dptr(x), dptr(s), int(g.var_tdim[v]), None)
    return x


def test_factor_jacobians_numeric(oracle, tiny):
    s, o = tiny
    g = s.graph
    base = o.vars
    rng = np.random.default_rng(0)
    for si, st in enumerate(g.stores):
        for k in rng.choice(st.count, min(3, st.count), replace=False):
            r0, Js, valid = o.eval_factor(si, int(k))
            assert valid
            for a, v in enumerate(st.vars[k]):
                d = g.var_tdim[v]; off = g.var_off[v]; n = G.DATA_DIM[int(g.var_kind[v])]
                eps = 1e-6
                num = np.zeros_like(Js[a])
                for c in range(d):
                    e = np.zeros(d); e[c] = eps
                    res = []
                    for sgn in (1, -1):
                        x = base.copy(); x[off:off + n] = _boxplus(oracle, g, v, base[off:off + n], sgn * e)
                        o.set(4, x)
                        res.append(o.eval_factor(si, int(k), jac=False)[0])
                    num[:, c] = (res[0] - res[1]) / (2 * eps)
                o.set(4, base)
                scale = max(1.0, np.abs(num).max())
                assert np.abs(num - Js[a]).max() < 2e-6 * scale, (st.type, a, np.abs(num - Js[a]).max())


def test_gradient_matches_cost_differences(oracle, tiny):
    s, o = tiny
    base = o.vars
    o.eval_grad_hess()
    grad = o.grad
    rng = np.random.default_rng(1)
    d = rng.normal(0, 1, o.order); d /= np.linalg.norm(d)
    eps = 1e-6
    cs = []
    for sgn in (1, -1):
        o.set(4, base); o.set(1, sgn * eps * d); o.apply_step(0)
        cs.append(o.eval_cost(False)[0])
    o.set(4, base)
    num = (cs[0] - cs[1]) / (2 * eps)
    assert abs(num - grad @ d) < 1e-5 * max(1.0, abs(num))


def test_hessian_is_jtj_and_solve_matches_dense(oracle, tiny):
    s, o = tiny
    o.eval_grad_hess()
    H = o.dense_hessian()
    assert np.allclose(H, H.T)
    assert np.linalg.eigvalsh(H).min() > -1e-6 * np.abs(H).max()
    o.damp(1e-3)
    Hd = o.dense_hessian(); g = o.grad
    o.factor(); o.solve()
    x = np.linalg.solve(Hd, g)
    assert np.abs(Hd @ o.step - g).max() < 1e-8 * np.abs(g).max()
    assert np.abs(x - o.step).max() < 1e-6 * np.abs(x).max()


def test_block_structure_rules(oracle, tiny):
    s, o = tiny
    rp, ci, bo, ps = o.structure()
    # rows sorted, explicit diagonal last (Optimizer.cpp:186-197); landmarks couple only to non-landmarks
    for i in range(o.nblocks):
        cols = ci[rp[i]:rp[i + 1]]
        assert cols[-1] == i and np.all(np.diff(cols) > 0)
        if i < s.n_points:
            assert len(cols) == 1
    assert ps[-1] == o.order


def test_lm_converges_and_is_deterministic(oracle, tiny):
    s, _ = tiny
    st = G.default_settings(40)
    a = Oracle(oracle, s.graph); b = Oracle(oracle, s.graph)
    sa, tpa, tna = a.optimize(st); sb, _, _ = b.optimize(st)
    assert sa.final_cost < 0.1 * sa.initial_cost and sa.stop_reason == 0
    assert sa.final_cost == sb.final_cost and np.array_equal(a.vars, b.vars)
    assert np.all(np.diff(tpa) <= 1e-9)


def test_parallel_baseline_mode_matches_serial(oracle):
    """bench.py's CPU-baseline mode (OpenMP factor loops, atomic adds, OpenBLAS band Cholesky) computes
    the same iteration as the serial checker path up to summation order."""
    from visual_inertial_bundle_adjustment_b200 import synth
    s = synth.make_session(**synth.CONFIGS["small_rs"])
    a, b = Oracle(oracle, s.graph), Oracle(oracle, s.graph)
    kind = b.set_parallel(4)
    _, a0, a1 = a.lm_iteration_timed(1e-5)
    _, b0, b1 = b.lm_iteration_timed(1e-5)
    assert abs(a0 - b0) <= 1e-12 * abs(a0) and abs(a1 - b1) <= 1e-9 * abs(a1)
    assert np.abs(a.step - b.step).max() <= 1e-7 * np.abs(a.step).max()
    assert kind in ("openblas-dpbtrf", "envelope")
    b.set_parallel(1)
    _, c0, c1 = b.lm_iteration_timed(1e-5)
    assert c0 == a0
