"""Covariance extraction (second half): Optimizer::computeJointCovariances / computeCovariances
(lib/small_thing/Optimizer.cpp:503-696).  The reference pins it with identities (tests/TestOptimizer.cpp:52-83): the
covariance of a set of variables is the inverse of their marginal information, H * Sigma = I to 1e-9.  Here: the block
of the dense inverse of the oracle's damped Hessian, and marginal-information * covariance = I."""
import numpy as np
import pytest
from tests.oracle_lib import Oracle
from visual_inertial_bundle_adjustment_b200 import synth, graph as G


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", ["tiny", "tiny_rs"])
def test_joint_covariance_is_a_block_of_the_inverse(oracle, cfg):
    from visual_inertial_bundle_adjustment_b200 import capi
    s = synth.make_session(**synth.CONFIGS[cfg])
    g = s.graph
    o = Oracle(oracle, g)
    o.eval_grad_hess()
    H = o.dense_hessian()
    lam = 1e-5
    Hd = H.copy()
    Hd[np.diag_indices_from(Hd)] = np.diag(H) * (1.0 + lam) + lam      # Optimizer::addDamping, Optimizer.cpp:136-146
    Sigma = np.linalg.inv(Hd)
    _, _, _, ps = o.structure()
    ctx = capi.Context(0)
    p = capi.Problem(ctx, g)
    rng = np.random.default_rng(1)
    cams = np.arange(g.num_points, o.nblocks)
    lists = [list(rng.choice(cams, 3, replace=False)), [int(cams[-1])], [0, int(cams[0]), 5], list(rng.choice(cams, 6, replace=False))]
    for idx in lists:
        cov, off, used = p.joint_covariance(idx, lam)
        assert used == lam
        rows = np.concatenate([np.arange(ps[i], ps[i + 1]) for i in idx])
        ref = Sigma[np.ix_(rows, rows)]
        assert off[-1] == rows.size and np.array_equal(np.diff(off), [ps[i + 1] - ps[i] for i in idx])
        assert np.abs(cov - ref).max() <= 1e-7 * np.abs(ref).max(), (idx, np.abs(cov - ref).max(), np.abs(ref).max())
        assert np.abs(cov - cov.T).max() <= 1e-9 * np.abs(cov).max()
        # the reference's identity: marginal information (Schur complement of the rest) times covariance = I
        rest = np.setdiff1d(np.arange(o.order), rows)
        Hm = Hd[np.ix_(rows, rows)] - Hd[np.ix_(rows, rest)] @ np.linalg.solve(Hd[np.ix_(rest, rest)], Hd[np.ix_(rest, rows)])
        # (Hm itself comes out of a numpy solve with the ill-conditioned rest block: entries ~1e4 against a covariance
        # ~1e-5, so the product carries ~1e-6 of rounding; the element-wise gate above is the parity gate)
        assert np.linalg.norm(Hm @ cov - np.eye(rows.size)) <= 1e-5
    # per-variable covariances (computeCovariances) are the diagonal blocks of a joint one
    cov3, off3, _ = p.joint_covariance(lists[0], lam)
    for k, i in enumerate(lists[0]):
        c1, _, _ = p.joint_covariance([i], lam)
        assert np.abs(c1 - cov3[off3[k]:off3[k + 1], off3[k]:off3[k + 1]]).max() <= 1e-9 * np.abs(c1).max()
    with pytest.raises(Exception, match="duplicated"):
        p.joint_covariance([int(cams[0]), int(cams[0])], lam)
    # all lists in one call (one factorisation), as SingleSessionProblem::computeCovariances does: the same blocks
    covs, used = p.joint_covariances(lists, lam)
    assert used == lam and len(covs) == len(lists)
    for idx, cv in zip(lists, covs):
        one, _, _ = p.joint_covariance(idx, lam)
        assert cv.shape == one.shape and np.abs(cv - one).max() <= 1e-9 * np.abs(one).max()
    with pytest.raises(Exception, match="duplicated"):
        p.joint_covariances([lists[0], [int(cams[0]), int(cams[0])]], lam)
    p.close(); ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", ["tiny", "tiny_rs"])
def test_marginal_problem_is_the_schur_complement(oracle, cfg):
    """Optimizer::computeMarginalProblem (Optimizer.cpp:422-494): H, b, cost against the dense Schur complement of the oracle's
    damped system onto the listed variables (the two routes agree to ~1e-9 in numpy; gates 1e-6), and the identity the
    reference tests, H * covariance = I (TestOptimizer.cpp:74-82)"""
    from visual_inertial_bundle_adjustment_b200 import capi
    s = synth.make_session(**synth.CONFIGS[cfg])
    g = s.graph
    o = Oracle(oracle, g)
    cost = o.eval_grad_hess()
    Hfull, grad = o.dense_hessian(), o.grad
    lam = 1e-5
    Hd = Hfull.copy()
    Hd[np.diag_indices_from(Hd)] = np.diag(Hfull) * (1.0 + lam) + lam
    _, _, _, ps = o.structure()
    ctx = capi.Context(0)
    p = capi.Problem(ctx, g)
    try:
        rng = np.random.default_rng(1)
        cams = np.arange(g.num_points, o.nblocks)
        for idx in ([int(cams[-1])], [0, int(cams[0]), 5], list(rng.choice(cams, 4, replace=False))):
            H, b, c, off, used = p.marginal_problem(idx, lam)
            assert used == lam
            rows = np.concatenate([np.arange(ps[i], ps[i + 1]) for i in idx])
            rest = np.setdiff1d(np.arange(o.order), rows)
            assert off[-1] == rows.size and np.array_equal(np.diff(off), [ps[i + 1] - ps[i] for i in idx])
            Hrr = Hd[np.ix_(rest, rest)]
            Hm = Hd[np.ix_(rows, rows)] - Hd[np.ix_(rows, rest)] @ np.linalg.solve(Hrr, Hd[np.ix_(rest, rows)])
            bm = grad[rows] - Hd[np.ix_(rows, rest)] @ np.linalg.solve(Hrr, grad[rest])
            cm = cost - 0.5 * grad[rest] @ np.linalg.solve(Hrr, grad[rest])
            assert np.linalg.norm(H - Hm) <= 1e-6 * np.linalg.norm(Hm), (idx, np.linalg.norm(H - Hm) / np.linalg.norm(Hm))
            assert np.linalg.norm(b - bm) <= 1e-6 * np.linalg.norm(bm), (idx, np.linalg.norm(b - bm) / np.linalg.norm(bm))
            assert abs(c - cm) <= 1e-9 * abs(cm), (c, cm)
            assert np.abs(H - H.T).max() <= 1e-9 * np.abs(H).max()
            cov, _, _ = p.joint_covariance(idx, lam)
            assert np.linalg.norm(H @ cov - np.eye(rows.size)) <= 1e-4
        with pytest.raises(Exception, match="duplicated"):
            p.marginal_problem([int(cams[0]), int(cams[0])], lam)
    finally:
        p.close(); ctx.close()
