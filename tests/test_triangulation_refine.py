"""Triangulation refinement (second half): SingleSessionAdapter::refineTriangulationResult,
viba/single_session/Triangulation.cpp:99-161 — undamped Gauss-Newton on every landmark with HuberLoss on the weighted
reprojection error, inlier classification on the unweighted image error, optional skipping of the outliers — run as
triangulatePoint runs it (:219-238): pass 1 (3.0 px, keep outliers, 3 iterations, radius 1.5), pass 2 (2.5 px, skip
outliers, 3 iterations, radius 1.0), Triangulation.h:29-39.

CPU: the oracle restatement against an independent numpy evaluation built from the oracle's per-factor residuals and
Jacobians (classification, one Gauss-Newton step, Huber weights), and its fixed point.
GPU: vb_refine_triangulation against the oracle on perturbed landmarks of global-shutter, rolling-shutter and
loop-closure sessions: inlier counts, inlier flags and status exactly, costs 1e-9 relative, refined points 1e-9 (a
1e-15 relative perturbation of the inputs moves them by < 1e-12 on these sessions)."""
import numpy as np
import pytest
from tests.oracle_lib import Oracle
from visual_inertial_bundle_adjustment_b200 import synth, graph as G

PASS1 = (3.0, False, 3, 1.5)
PASS2 = (2.5, True, 3, 1.0)


def perturbed_session(name, sigma=0.05, seed=5):
    s = synth.make_session(**synth.CONFIGS[name])
    g = s.graph
    rng = np.random.default_rng(seed)
    for v in np.asarray(s.ids["point"]):
        g.var_data[g.var_off[v]:g.var_off[v] + 3] += rng.normal(0, sigma, 3)
    return s


def visual_stores(g):
    return [(si, st) for si, st in enumerate(g.stores) if st.type in (G.F_VISUAL, G.F_VISUAL_RS)]


def numpy_one_iteration(o, g, threshold, skip, radius):
    """numInliers, flags and the Gauss-Newton step of one iteration, landmark by landmark, from orc_eval_factor"""
    H, grad, n_in, flags = {}, {}, {}, []
    for si, st in visual_stores(g):
        f = np.zeros(st.vars.shape[0], np.uint8)
        for k in range(st.vars.shape[0]):
            v = int(st.vars[k, 0])
            H.setdefault(v, np.zeros((3, 3))); grad.setdefault(v, np.zeros(3)); n_in.setdefault(v, 0)
            r, Js, valid = o.eval_factor(si, k)
            if not valid:
                continue
            sqrt_h = st.consts[k, 2:6].reshape(2, 2)
            e = np.linalg.solve(sqrt_h, r)
            if e @ e < threshold * threshold:
                n_in[v] += 1; f[k] = 1
            elif skip:
                continue
            s2 = r @ r
            d = radius / np.sqrt(s2) if s2 > radius * radius else 1.0
            J = Js[0]
            grad[v] += d * (r @ J); H[v] += d * (J.T @ J)
        flags.append(f)
    return H, grad, n_in, np.concatenate(flags)


@pytest.mark.parametrize("name,settings", [("tiny", PASS1), ("tiny", PASS2), ("tiny_rs", PASS2)])
def test_oracle_iteration_against_numpy(oracle, name, settings):
    s = perturbed_session(name)
    g = s.graph
    thr, skip, _, radius = settings
    o = Oracle(oracle, g)
    if settings is PASS2:      # as triangulatePoint does: the second pass starts from the result of the first
        o.refine_triangulation(*PASS1)
    H, grad, n_in, flags = numpy_one_iteration(o, g, thr, skip, radius)
    x0 = o.vars.copy()
    pp, fl, fails = o.refine_triangulation(thr, skip, 1, radius)
    assert np.array_equal(fl, flags)
    lms = sorted(H)
    assert pp.shape[0] == len(lms)
    assert 0 < flags.sum() < flags.size, "the scenario must hold inliers and outliers"
    x1 = o.vars
    n_failed = 0
    for j, v in enumerate(lms):
        assert pp[j, 0] == n_in[v]
        off = g.var_off[v]
        if n_in[v] == 0 and skip:     # nothing entered the normal equations: not positive definite, the point is kept
            assert pp[j, 2] == -1 and np.array_equal(x0[off:off + 3], x1[off:off + 3])
            n_failed += 1
            continue
        assert pp[j, 2] == 0
        step = np.linalg.solve(H[v], grad[v])
        assert np.abs((x0[off:off + 3] - step) - x1[off:off + 3]).max() <= 1e-9 * max(1.0, np.abs(step).max())
    assert fails == n_failed
    other = np.ones(x0.size, bool)
    for v in lms:
        other[g.var_off[v]:g.var_off[v] + 3] = False
    assert np.array_equal(x0[other], x1[other]), "only the landmarks move"


def test_oracle_fixed_point_and_failure(oracle):
    s = perturbed_session("tiny")
    g = s.graph
    o = Oracle(oracle, g)
    o.refine_triangulation(3.0, False, 60, 1.5)        # Gauss-Newton on a Huber cost converges linearly
    x = o.vars.copy()
    pp, _, fails = o.refine_triangulation(3.0, False, 1, 1.5)
    assert fails == 0 and np.abs(o.vars - x).max() <= 1e-8      # converged: another step does not move the points
    # a landmark pushed far behind everything: whatever the cameras make of it, the bookkeeping must be consistent
    v = int(np.asarray(s.ids["point"])[0])
    far = o.vars.copy()
    far[g.var_off[v]:g.var_off[v] + 3] = [0.0, 0.0, -1e7]
    g.var_data[:] = far
    o2 = Oracle(oracle, g)
    pp2, fl2, fails2 = o2.refine_triangulation(2.5, True, 3, 1.0)
    lms = sorted(set(int(a) for _, st in visual_stores(g) for a in st.vars[:, 0]))
    j = lms.index(v)
    assert fails2 == int((pp2[:, 2] == -1).sum())
    if pp2[j, 2] == -1:   # empty normal equations (every projection failed or was an outlier): the point stays
        assert pp2[j, 0] == 0
        assert np.array_equal(o2.vars[g.var_off[v]:g.var_off[v] + 3], far[g.var_off[v]:g.var_off[v] + 3])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["tiny", "small_rs", "small_lc_rs"])
def test_device_triangulation_refinement_matches_oracle(oracle, name):
    from visual_inertial_bundle_adjustment_b200 import capi
    s = perturbed_session(name)
    g = s.graph
    o = Oracle(oracle, g)
    ctx = capi.Context(0)
    p = capi.Problem(ctx, g)
    try:
        moved = 0.0
        x_start = o.vars.copy()
        for settings in (PASS1, PASS2):
            pp_o, fl_o, fails_o = o.refine_triangulation(*settings)
            pp_g, fl_g, fails_g = p.refine_triangulation(*settings)
            assert fails_g == fails_o == 0
            assert np.array_equal(pp_g[:, 0], pp_o[:, 0]), "numInlierObservations"
            assert np.array_equal(pp_g[:, 2], pp_o[:, 2]), "status"
            assert np.array_equal(fl_g, fl_o), "inlierObservationIndices"
            assert np.abs(pp_g[:, 1] - pp_o[:, 1]).max() <= 1e-9 * pp_o[:, 1].max()
            xo, xg = o.vars, p.vars
            assert np.abs(xg - xo).max() <= 1e-9
            moved = max(moved, np.abs(xo - x_start).max())
            assert 0 < fl_o.sum() < fl_o.size
        assert moved > 1e-3, "the refinement must have moved the landmarks"
        # the refined state is a regular state of the problem: the cost of the whole graph agrees too
        c_o = o.eval_cost(False)[0]
        c_g = p.compute_cost(False)[0]
        assert abs(c_g - c_o) <= 1e-8 * abs(c_o)
    finally:
        p.close(); ctx.close()
