"""Committed golden vectors (tests/golden/*.npz, written by tests/golden/make_golden.py from the CPU
oracle; see that script's header for what they can and cannot pin).

* no GPU: the session generator still produces the same graph (digest) and the oracle still produces the
  same numbers — cost, gradient, per-factor residuals and Jacobians, damped step, a whole LM run;
* GPU: the CUDA path through the C ABI against the same fixed numbers, with nothing from oracle/ loaded:
  cost rel 1e-12, gradient 1e-10 of scale, step 1e-7, per-run final cost rel 1e-9, final state 1e-6."""
import os
import sys

import numpy as np
import pytest

from visual_inertial_bundle_adjustment_b200 import synth, capi, graph as G

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden as MG  # noqa: E402


def load(name):
    return np.load(os.path.join(HERE, "golden", name + ".npz"))


@pytest.mark.parametrize("name", MG.SESSIONS)
def test_oracle_reproduces_golden(name, oracle):
    z = load(name)
    d = MG.golden_for(name, oracle)
    assert str(d["digest"]) == str(z["digest"]), "the synthetic session changed: regenerate tests/golden deliberately"
    assert set(d) == set(z.files)
    for k in z.files:
        if k in ("digest", "imu_digest"):
            assert str(d[k]) == str(z[k]), k
            continue
        a, b = np.asarray(d[k], np.float64), np.asarray(z[k], np.float64)
        assert a.shape == b.shape, k
        assert np.abs(a - b).max() <= 1e-11 * max(1.0, np.abs(b).max()), k


def to_registration_order(g, p, vec):
    """permuted device vector -> registration order (Variable::index order), from the graph alone"""
    reg = g.var_index >= 0
    nblocks = int(g.var_index.max()) + 1
    tdim = np.zeros(nblocks, np.int64)
    tdim[g.var_index[reg]] = g.var_tdim[reg]
    ps_reg = np.concatenate([[0], np.cumsum(tdim)])
    ps_dev = p.param_start()
    out = np.zeros(int(ps_reg[-1]))
    for b in range(nblocks):
        out[ps_reg[b]:ps_reg[b + 1]] = vec[ps_dev[b]:ps_dev[b] + tdim[b]]
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("name", MG.SESSIONS)
def test_cuda_path_matches_golden(name):
    z = load(name)
    s = synth.make_session(**synth.CONFIGS[name])
    g = s.graph
    assert MG.graph_digest(g) == str(z["digest"])
    ctx = capi.Context(0)
    p = capi.Problem(ctx, g)
    assert p.order == int(z["order"]) and p.num_blocks == int(z["nblocks"])
    c0 = p.compute_grad_hess()
    assert abs(c0 - float(z["cost0"])) <= 1e-12 * abs(float(z["cost0"]))
    grad = to_registration_order(g, p, p.grad)
    assert np.abs(grad - z["grad"]).max() <= 1e-10 * np.abs(z["grad"]).max()
    c, st = p.compute_cost(True)
    assert abs(c - float(z["cost_eval"])) <= 1e-12 * abs(float(z["cost_eval"])) and tuple(st) == tuple(int(x) for x in z["cost_stats"])
    p.add_damping(MG.LAMBDA); p.step_from_grad(); p.factor(); p.solve(0)
    step = to_registration_order(g, p, p.step)
    assert np.abs(step - z["step"]).max() <= 1e-7 * np.abs(z["step"]).max()
    assert abs(0.5 * p.dot(0, 2) - float(z["model_decrease"])) <= 1e-9 * abs(float(z["model_decrease"]))
    p.close()
    p = capi.Problem(ctx, g)
    summ = p.optimize(G.default_settings(MG.ITERS))
    assert summ.num_iterations == int(z["opt_iterations"])
    assert abs(summ.initial_cost - float(z["opt_initial_cost"])) <= 1e-12 * abs(float(z["opt_initial_cost"]))
    assert abs(summ.final_cost - float(z["opt_final_cost"])) <= 1e-9 * abs(float(z["opt_final_cost"]))
    assert np.abs(p.vars - z["opt_vars"]).max() <= 1e-6
    p.close(); ctx.close()
