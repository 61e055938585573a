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
    p.close(); ctx.close()
