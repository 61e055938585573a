#!/usr/bin/env python
"""BASELINE config C5 — factorisation-only sweep: landmark Schur complement + supernodal Cholesky
(FP64 DMMA) + solves on reduced camera systems of growing pose counts, beside the CPU stand-in for
the reference's BaSpaCho/OpenBLAS path (oracle baseline mode: OpenMP landmark elimination + LAPACK
dpbtrf/dpbtrs on scipy's OpenBLAS, all host cores).

  python profiles/c5_factor_sweep.py [--poses 5000 10000 20000] [--obs-per-rig 500] [--steps 5] [--cpu-max-poses 10000]

Sessions are synthetic Aria-like walks at 10 Hz (poses = duration x 10) with a sparse landmark field so
that generation stays cheap; the Hessian values are the real assembled J^T J (one vb_eval_grad_hess
+ damping per timed factorisation, outside the timed region: the factorisation is in place).
Prints one JSON line per size.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--poses", type=int, nargs="+", default=[5000, 10000, 20000])
    ap.add_argument("--obs-per-rig", type=int, default=500)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--cpu-max-poses", type=int, default=10000)
    a = ap.parse_args()
    import torch
    from visual_inertial_bundle_adjustment_b200 import capi, synth
    from tests.oracle_lib import load_oracle, Oracle
    ctx = capi.Context(0)
    peak = ctx.dmma_peak_tflops()
    for n in a.poses:
        dur = n / 10.0
        s = synth.make_session(duration_s=dur, n_landmarks=int(n * a.obs_per_rig / 50), n_obs=n * a.obs_per_rig)
        p = capi.Problem(ctx, s.graph)
        w = p.work()
        fac, sol = [], []
        for it in range(a.steps + 2):
            p.compute_grad_hess()
            p.add_damping(1e-5)
            p.step_from_grad()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            p.factor()      # synchronises (numeric check)
            t1 = time.perf_counter()
            p.solve(0)      # synchronises
            t2 = time.perf_counter()
            if it >= 2:
                fac.append((t1 - t0) * 1e3)
                sol.append((t2 - t1) * 1e3)
        line = {"poses": n, "observations": int(s.n_obs), "landmarks": int(s.n_points), "order": int(p.order),
                "hess_doubles": int(p.data_size), "lumps": int(p.num_lumps), "levels": int(p.num_levels),
                "schur_gflop": w["schur_flops"] / 1e9, "chol_gflop": w["chol_flops"] / 1e9,
                "gpu_factor_ms": float(np.mean(fac)), "gpu_solve_ms": float(np.mean(sol)),
                "gpu_tflops": (w["schur_flops"] + w["chol_flops"]) / (np.mean(fac) * 1e-3) / 1e12,
                "dmma_peak_tflops": peak, "timing": "host wall clock around vb_factor / vb_solve (each synchronises)"}
        line["frac_of_dmma_peak"] = line["gpu_tflops"] / peak
        p.close()
        if n <= a.cpu_max_poses:
            o = Oracle(load_oracle(), s.graph)
            cores = os.cpu_count() or 1
            kind = o.set_parallel(cores)
            ms = [o.lm_iteration_timed(1e-5)[0] for _ in range(2)][-1]
            line.update({"cpu_factor_ms": float(ms[2]), "cpu_solve_ms": float(ms[3]), "cpu_cores": cores, "cpu_solver": kind,
                         "speedup_factor": float(ms[2] / np.mean(fac))})
            del o
        print(json.dumps(line), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
