"""Measurement of the iterative path on one GPU: PCG on the reduced camera system with the
reference's default settings (pcgMaxIterations 40, pcgDesiredResidual 1e-10, Optimizer.h:43-45) and run to 1e-8, at a
named workload, beside the direct factor + solve on the same damped system.

    python profiles/pcg_bench.py [workload] > profiles/r02_pcg_<workload>.json

Preconditioners: identity, block Jacobi, block Gauss-Seidel (pseudo-factor + two supernodal solves per application) and
lower-precision solve (complete factor of the float-rounded copy).  One PCG iteration = one symmetric matrix-vector product over the unfactored supernode panels (every camera-panel double
read once: HBM-bound, algorithmic bytes = 8 x camera-panel doubles) + preconditioner + 2 fused dot kernels + 2 axpy
kernels + one 48-byte read-back.  Wall-clock around calls that end in a stream synchronize; not under a profiler."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import bench  # noqa: E402
from visual_inertial_bundle_adjustment_b200 import capi, graph as G  # noqa: E402


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "c2_5min"
    lam = 1e-5
    s, _ = bench.make_workload(workload)
    ctx = capi.Context(0)
    p = capi.Problem(ctx, s.graph)
    bench.attach_rs_imu(p, s)
    cam_doubles = float(p.shard_info()["reduced_doubles"])
    peaks, _ = bench.read_peaks()

    def system():
        p.compute_grad_hess(); p.add_damping(lam); p.step_from_grad()

    def timed(fn):
        t = time.perf_counter(); fn()   # vb_factor / vb_solve return after a stream synchronize
        return (time.perf_counter() - t) * 1e3

    out = {"workload": workload, "lambda": lam, "order": int(p.order), "camera_panel_doubles": cam_doubles, "runs": []}
    # direct, for the comparison (second run: warm)
    for _ in range(2):
        p.set_linear_solver(G.SOLVER_DIRECT); system()
        tf = timed(p.factor); ts = timed(lambda: p.solve(0))
    direct = p.step.copy()
    out["direct"] = {"factor_ms": tf, "solve_ms": ts}
    plans = [(G.SOLVER_PCG_JACOBI, "jacobi", ((40, 1e-10), (2000, 1e-8))), (G.SOLVER_PCG_TRIVIAL, "trivial", ((40, 1e-10), (2000, 1e-8))),
             (G.SOLVER_PCG_GAUSS_SEIDEL, "gauss-seidel", ((40, 1e-10), (400, 1e-8))),
             (G.SOLVER_PCG_LOWER_PREC_SOLVE, "lower-prec-solve", ((40, 1e-10),))]
    for solver, label, settings in plans:
        for max_it, res in settings:
            for _ in range(2):
                p.set_linear_solver(solver, max_it, res); system()
                tf = timed(p.factor); ts = timed(lambda: p.solve(0))
            it, r = p.pcg_last()
            x = p.step
            per_it = ts / max(it, 1)
            out["runs"].append({"precond": label, "max_iterations": max_it, "desired_residual": res, "iterations": it, "residual": r,
                                "factor_ms": tf, "solve_ms": ts, "ms_per_pcg_iteration": per_it,
                                "matvec_algorithmic_gbs_lower_bound": 8 * cam_doubles / (per_it * 1e-3) / 1e9,
                                "hbm_peak_gbs": peaks.get("hbm_gbs"),
                                "step_rel_diff_to_direct": float(np.abs(x - direct).max() / np.abs(direct).max())})
    p.close(); ctx.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
