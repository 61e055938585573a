"""vb_dot_no_alloc — the allocation-free dot product (include/viba_b200.h; no counterpart in the reference, added
for callers that cannot call cudaMalloc).  Its scratch is a fixed __device__ array inside
the library, shared by every call on a GPU.

Collected last on purpose: it must pass alone in a fresh process and after every other
test.  The GPU test calls the new entry point through the C ABI on device copies of a problem's step / gradient vectors
and checks it against the existing entry point (vb_dot, whose accumulator is cudaMalloc'ed scratch) on two calls in a
row with different inputs and lengths — so a stale ticket or stale partial sums of the previous call would show — at
1e-12 of sum |a_i b_i| (the cost gate of tests/test_gpu_parity.py; the two kernels sum in different orders)."""
import numpy as np
import pytest
from visual_inertial_bundle_adjustment_b200 import capi, synth


def test_sizes_beyond_the_fixed_scratch_are_refused():
    """argument checks come before any device work, so this runs without a GPU too"""
    for n in (capi.DOT_NO_ALLOC_MAX_N + 1, 1 << 40):
        with pytest.raises(capi.VbError) as e:
            capi.dot_no_alloc(0, 256, 256, n, 256)
        assert e.value.code == 1 and "exceeds VB_DOT_NO_ALLOC_MAX_N" in str(e.value)
    with pytest.raises(capi.VbError) as e:
        capi.dot_no_alloc(0, 256, 256, -1, 256)
    assert e.value.code == 1
    with pytest.raises(capi.VbError) as e:
        capi.dot_no_alloc(0, 0, 256, 8, 256)
    assert e.value.code == 1 and "null" in str(e.value)


@pytest.mark.gpu
def test_no_alloc_dot_matches_vb_dot_on_two_calls_in_a_row():
    import torch
    s = synth.make_session(**synth.CONFIGS["c1_60s"])     # order ~ 66 000: several CTAs, so the ticket path runs
    ctx = capi.Context(0)
    p = capi.Problem(ctx, s.graph)
    try:
        p.compute_grad_hess(); p.add_damping(1e-4); p.step_from_grad(); p.factor(); p.solve(0)
        step, grad = p.step, p.grad
        ref_sg, ref_gg, ref_ss = p.dot(0, 2), p.dot(2, 2), p.dot(0, 0)     # the existing entry point
    finally:
        p.close(); ctx.close()
    n = step.size
    assert n > 3 * 16384
    dev = torch.device("cuda:0")
    a = torch.from_numpy(step).to(dev); b = torch.from_numpy(grad).to(dev)
    out = torch.full((4,), float("nan"), dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    torch.cuda.synchronize(dev)
    before = torch.cuda.memory_allocated(dev)

    def call(x, y, length, slot):
        capi.dot_no_alloc(0, x.data_ptr(), y.data_ptr(), length, out.data_ptr() + 8 * slot, stream)

    # two calls in a row on one stream, no synchronisation between them, different inputs
    call(a, b, n, 0)
    call(b, b, n, 1)
    # a shorter vector (one CTA: partial sums of the previous call stay behind in the scratch), then the first again
    m = 5000
    call(a, a, m, 2)
    call(a, b, n, 3)
    torch.cuda.synchronize(dev)
    assert torch.cuda.memory_allocated(dev) == before
    got = out.cpu().numpy()

    def tol(x, y):
        return 1e-12 * float(np.abs(x * y).sum())
    assert abs(got[0] - ref_sg) <= tol(step, grad), (got[0], ref_sg)
    assert abs(got[1] - ref_gg) <= tol(grad, grad), (got[1], ref_gg)
    assert abs(got[2] - float(step[:m] @ step[:m])) <= tol(step[:m], step[:m])
    assert got[3] == got[0], "fixed summation order: the same call gives the same bits"
    # whole-vector variant of the short call against the existing entry point as well
    call(a, a, n, 2)
    call(a, a, 0, 1)
    torch.cuda.synchronize(dev)
    got = out.cpu().numpy()
    assert abs(got[2] - ref_ss) <= tol(step, step), (got[2], ref_ss)
    assert got[1] == 0.0
