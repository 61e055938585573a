#!/usr/bin/env python
"""Regenerates tests/golden/*.npz.

What these fixtures are — and are not.  The reference's own implementation of the path cannot be
compiled or imported in this environment (its BaSpaCho / Eigen / Sophus / dispenso dependencies are
not vendored and there is no network: SURVEY.md §8c), and its tests hold no golden files for this
path.  The vectors here are therefore outputs of the CPU oracle (oracle/, the restatement of the
reference's algorithm that is itself pinned by the reference's test identities in
tests/test_oracle_math.py / test_oracle_engine.py) on seeded synthetic sessions.  They pin
(a) the oracle and the session generator against drift (tests/test_golden.py, no GPU) and
(b) the CUDA path against fixed numbers without the oracle in the loop (tests/test_golden.py -m gpu).

  python tests/golden/make_golden.py          # rewrites the .npz files next to this script
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from visual_inertial_bundle_adjustment_b200 import synth, graph as G  # noqa: E402
from tests.oracle_lib import load_oracle, Oracle  # noqa: E402

SESSIONS = ["tiny", "tiny_rs", "tiny_lc"]
LAMBDA = 1e-4
ITERS = 8
FACTORS_PER_STORE = 3


def graph_digest(g):
    h = hashlib.sha256()
    for a in (g.var_kind, g.var_tdim, g.var_index, g.var_data):
        h.update(np.ascontiguousarray(a).tobytes())
    for st in g.stores:
        h.update(st.vars.tobytes()); h.update(st.consts.tobytes())
        if st.aux is not None:
            h.update(st.aux.tobytes())
    return h.hexdigest()


def golden_for(name, lib):
    s = synth.make_session(**synth.CONFIGS[name])
    g = s.graph
    o = Oracle(lib, g)
    out = {"digest": np.array(graph_digest(g)), "order": np.int64(o.order), "nblocks": np.int64(o.nblocks)}
    out["cost0"] = np.float64(o.eval_grad_hess())
    out["grad"] = o.grad
    c, st = o.eval_cost(True)
    out["cost_eval"], out["cost_stats"] = np.float64(c), np.array(st, np.int64)
    res, jac = [], []
    for si, store in enumerate(g.stores):
        for k in range(min(FACTORS_PER_STORE, store.count)):
            r, Js, valid = o.eval_factor(si, k)
            res.append(np.concatenate([[si, k, float(valid)], r]))
            jac.append(np.concatenate([J.ravel() for J in Js]))
    out["factor_residuals"] = np.concatenate(res)
    out["factor_jacobians"] = np.concatenate(jac)
    o.damp(LAMBDA); o.factor(); o.solve()
    out["step"] = o.step
    out["model_decrease"] = np.float64(0.5 * o.step @ o.grad)
    if any(st.type == G.F_VISUAL_RS for st in g.stores):
        # rolling-shutter tables recomputed from raw IMU samples with the session's initial calibration
        # (RollingShutterData::compute, oracle/rs_refresh.hpp) and the cost evaluated on them
        d = synth.imu0_samples(s)
        o3 = Oracle(lib, g)
        o3.rs_set_imu(d["store"], d["time_ns"], d["gyro"], d["accel"], d["rig_mid_time_us"], d["rig_calib_var"], d["half_length_us"])
        o3.rs_refresh()
        rigs = np.flatnonzero(d["rig_calib_var"] >= 0)
        out["rs_tables_refreshed"] = o3.store_aux(d["store"])[4:].reshape(-1, synth.RS_STRIDE)[rigs]
        out["rs_cost_after_refresh"] = np.float64(o3.eval_cost(False)[0])
        out["imu_digest"] = np.array(hashlib.sha256(d["time_ns"].tobytes() + d["gyro"].tobytes() + d["accel"].tobytes()).hexdigest())
    o2 = Oracle(lib, g)
    summ, tp, tn = o2.optimize(G.default_settings(ITERS))
    out["opt_iterations"] = np.int64(summ.num_iterations)
    out["opt_initial_cost"], out["opt_final_cost"] = np.float64(summ.initial_cost), np.float64(summ.final_cost)
    out["opt_trace_prev"], out["opt_trace_new"] = tp, tn
    out["opt_vars"] = o2.vars
    return out


def main():
    lib = load_oracle()
    for name in SESSIONS:
        d = golden_for(name, lib)
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **d)
        print(name, os.path.getsize(path), "bytes", "order", int(d["order"]), "cost0", float(d["cost0"]), "->", float(d["opt_final_cost"]))


if __name__ == "__main__":
    main()
