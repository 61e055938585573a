"""Times vb_lm_iteration_timed on one workload under a list of environment settings of the symbolic phase
(VB_ND_LEAF_BLOCKS, VB_MAX_LUMP_WIDTH, VB_GRP_MAX_ROWS, VB_GRP_CHUNKS, ...): one JSON line per setting.
  python profiles/sweep_env.py c2_5min_rs "VB_ND_LEAF_BLOCKS=150" "VB_ND_LEAF_BLOCKS=600,VB_MAX_LUMP_WIDTH=128" ...
"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench
from visual_inertial_bundle_adjustment_b200 import capi


def main():
    workload, settings = sys.argv[1], sys.argv[2:] or [""]
    s, cfg = bench.make_workload(workload)
    ctx = capi.Context(0)
    print(json.dumps({"dmma_shapes_tflops": ctx.dmma_shapes_tflops()}), flush=True)
    for setting in settings:
        keys = []
        for kv in filter(None, setting.split(",")):
            k, v = kv.split("=")
            os.environ[k] = v
            keys.append(k)
        t0 = time.time()
        p = capi.Problem(ctx, s.graph)
        setup = time.time() - t0
        bench.attach_rs_imu(p, s)
        for _ in range(3):
            p.lm_iteration_timed(1e-5)
        acc = {}
        n = 5
        for _ in range(n):
            t, c0, c1 = p.lm_iteration_timed(1e-5)
            for k, v in t.items():
                acc[k] = acc.get(k, 0.0) + v / n
        w = p.work()
        print(json.dumps({"workload": workload, "setting": setting, "setup_s": round(setup, 1), "levels": p.num_levels, "lumps": p.num_lumps,
                          "hess_doubles": p.data_size, "schur_gflop": round(w["schur_flops"] / 1e9, 1), "chol_gflop": round(w["chol_flops"] / 1e9, 1),
                          "phases_ms": {k: round(v, 3) for k, v in acc.items() if k != "step_dot_grad"}, "cost": [c0, c1]}), flush=True)
        p.close()
        for k in keys:
            del os.environ[k]
    ctx.close()


if __name__ == "__main__":
    main()
