#!/usr/bin/env python
"""Timing of the rolling-shutter table refresher (K12) beside the CPU restatement:
vb_rs_refresh (launch + rsRefreshKernel + status read-back, host wall clock around the synchronising call)
against the oracle's RollingShutterData::compute loop on one host core and on all of them, for sessions
with 3 000 (C2) and 12 000 (C3) rigs.  The sessions carry few landmarks so that they build quickly; the
refresher's work depends only on the number of rolling-shutter rigs.  One JSON line per size.

  python profiles/rs_refresh_timing.py [--durations 300 1200] [--reps 20]
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
    ap.add_argument("--durations", type=float, nargs="+", default=[300.0, 1200.0])
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    from visual_inertial_bundle_adjustment_b200 import capi, synth
    from tests.oracle_lib import load_oracle, Oracle
    ctx = capi.Context(0)
    for dur in a.durations:
        s = synth.make_session(duration_s=dur, n_landmarks=int(dur * 20), n_obs=int(dur * 1000), rgb=True, n_imus=2)
        d = synth.imu0_samples(s)
        args = (d["store"], d["time_ns"], d["gyro"], d["accel"], d["rig_mid_time_us"], d["rig_calib_var"], d["half_length_us"])
        p = capi.Problem(ctx, s.graph)
        p.rs_set_imu(*args)
        for _ in range(3):
            p.rs_refresh()
        t0 = time.perf_counter()
        for _ in range(a.reps):
            p.rs_refresh()
        gpu_ms = (time.perf_counter() - t0) * 1e3 / a.reps
        tab_gpu = p.store_aux(d["store"])
        p.close()
        o = Oracle(load_oracle(), s.graph)
        o.rs_set_imu(*args)
        o.rs_refresh()
        t0 = time.perf_counter()
        for _ in range(3):
            o.rs_refresh()
        cpu1_ms = (time.perf_counter() - t0) * 1e3 / 3
        cores = os.cpu_count() or 1
        o.set_parallel(cores)
        o.rs_refresh()
        t0 = time.perf_counter()
        for _ in range(3):
            o.rs_refresh()
        cpuN_ms = (time.perf_counter() - t0) * 1e3 / 3
        err = float(np.abs(tab_gpu - o.store_aux(d["store"])).max())
        rigs = int((d["rig_calib_var"] >= 0).sum())
        print(json.dumps({"rigs": int(s.n_rigs), "rs_rigs": rigs, "imu_samples": int(d["time_ns"].size),
                          "gpu_ms_per_refresh": gpu_ms, "cpu_ms_1_core": cpu1_ms, "cpu_ms_all_cores": cpuN_ms, "cores": cores,
                          "table_bytes": rigs * synth.RS_STRIDE * 8, "max_abs_diff_vs_oracle": err,
                          "timing": "host wall clock around the synchronising vb_rs_refresh call"}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
