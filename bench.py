#!/usr/bin/env python
"""bench.py — ms per LM iteration (Jacobian+Hessian build, Schur + Cholesky solve) on synthetic
Aria-like sessions, the metric of BASELINE.json.  One "step" = one accepted-path LM iteration
(Optimizer.cpp:800-1076): grad/Hess build, damping, factor, solve, backup+apply, cost.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

N>1 is launched by torchrun (one rank per GPU): every rank holds the SAME session, landmarks (with
all of their observations), factors and the subtrees of the reduced system's elimination tree are
split over the ranks; the shared top of the reduced system, the camera right-hand side / step and the
landmark steps are summed over NCCL at the exchange points of the C ABI's collective hook.  Total work is fixed as N grows: "scaling": "strong".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

DEFAULT_WORKLOAD = "c2_5min_rs"
SAMPLE_WORKLOAD = "c1_60s_rs"  # bounded CPU sample: 60 s of the same kind of session


def read_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [x.strip() for x in out.strip().split(",")]
                if len(parts) >= 6:
                    self.rows.append(parts)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons, "samples": len(sm)}


def make_workload(name, rank=0):
    from visual_inertial_bundle_adjustment_b200 import synth
    cfg = dict(synth.CONFIGS[name])
    cfg["seed"] = 42  # every rank builds the same session (rank is kept for callers that want replicas)
    return synth.make_session(**cfg), cfg


def attach_rs_imu(target, session):
    """raw IMU-0 samples for the rolling-shutter store, if the session has one; `target` is a capi.Problem or
    the oracle.  Returns a note for the JSON line; a failure leaves the tables of the session as they are."""
    from visual_inertial_bundle_adjustment_b200 import synth, graph as G
    if not any(st.type == G.F_VISUAL_RS for st in session.graph.stores):
        return "no rolling-shutter camera"
    d = synth.imu0_samples(session)
    try:
        target.rs_set_imu(d["store"], d["time_ns"], d["gyro"], d["accel"], d["rig_mid_time_us"], d["rig_calib_var"], d["half_length_us"])
        target.rs_refresh(d["store"])
        return "RollingShutterData::compute for %d rigs inside every iteration (grad_hess phase)" % int((d["rig_calib_var"] >= 0).sum())
    except Exception as ex:  # keep the bench alive: measured without the refresh, and say so
        if hasattr(target, "rs_clear_imu"):
            target.rs_clear_imu(d["store"])
        return "off (%s)" % str(ex)[:120]


def cpu_leg(workload, steps, warmup):
    """The reference's CPU path cannot be built offline (BaSpaCho/Eigen/Sophus absent: SURVEY.md
    F1,F3), so the CPU arm is the oracle port in its baseline mode — factor loops and landmark
    elimination on every host core (OpenMP, atomic block adds, as the reference's thread pool does
    with locked adds) and the reduced-system Cholesky through LAPACK dpbtrf on the multi-threaded
    OpenBLAS that ships with scipy (the reference: BaSpaCho's BLAS backend on OpenBLAS) — timed on
    a bounded sample (60 s of the same kind of session) and scaled by session duration."""
    from tests.oracle_lib import load_oracle, Oracle
    from visual_inertial_bundle_adjustment_b200 import synth
    sample = SAMPLE_WORKLOAD if synth.CONFIGS[workload].get("rgb") else "c1_60s"
    s, _ = make_workload(sample)
    o = Oracle(load_oracle(), s.graph)
    cores = os.cpu_count() or 1
    solver_kind = o.set_parallel(cores)
    attach_rs_imu(o, s)
    scale = synth.CONFIGS[workload]["duration_s"] / synth.CONFIGS[sample]["duration_s"]
    ts = []
    for i in range(warmup + steps):
        ms, c0, c1 = o.lm_iteration_timed(1e-5)
        if i >= warmup:
            ts.append(ms)
    ms = np.mean(np.array(ts), 0)
    return {"value": float(ms[6] * scale), "unit": "ms/iteration", "cores": cores, "kind": "port",
            "solver": solver_kind,
            "sample": f"oracle port ({cores} OpenMP threads, reduced system by {solver_kind}) on the {sample} session ({s.n_obs + s.ext.get("rs_obs", 0)} observations, {s.n_rigs} rigs), "
                      f"x{scale:g} by session duration to {workload}; split g/H {ms[0]:.0f} ms, factor {ms[2]:.0f} ms, "
                      f"solve {ms[3]:.0f} ms, cost {ms[5]:.0f} ms on the sample",
            "sample_ms": float(ms[6]), "scale": scale}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    metric = "ms per LM iteration (Jacobian+Hessian build, Schur+Cholesky solve)"
    config = {"workload": f"synthetic {a.workload} Aria Gen1-like session, the same session at every N (2 global-shutter SLAM cameras + RGB rolling-shutter "
                          "camera with readout/time-offset states, 2 IMUs, 5 s calibration windows, RW + factory priors)",
              "l2": "inputs larger than L2 (Hessian + factor tables exceed 126 MB); no flush needed"}

    if a.impl == "reference":
        if rank != 0:
            return 0
        k, w = min(a.steps, 3), min(a.warmup, 1)
        cb = cpu_leg(a.workload, k, w)
        line = {"impl": "reference", "metric": metric, "value": cb["value"], "unit": "ms/iteration", "n_gpus": a.gpus,
                "steps": k, "warmup": w, "ms_per_step": cb["value"], "higher_is_better": False, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config, "cpu_baseline": cb,
                "e2e": {"value": cb["value"], "unit": "ms/iteration", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return 0

    # stdout carries exactly ONE JSON line: library chatter on fd 1 (e.g. NCCL's version banner) goes to stderr
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    from visual_inertial_bundle_adjustment_b200 import capi
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_
        torch.cuda.set_device(local_rank)
        dist_.init_process_group("nccl")
        dist = dist_
    s, cfg = make_workload(a.workload, rank)
    t0 = time.time()
    coll = None
    if world > 1:
        from visual_inertial_bundle_adjustment_b200 import sharding
        ctx, p, coll = sharding.make_sharded_problem(s.graph, rank, world, local_rank)
    else:
        ctx = capi.Context(local_rank)
        p = capi.Problem(ctx, s.graph)
    setup_s = time.time() - t0
    lam = 1e-5
    # the reference refreshes the rolling-shutter tables from the raw IMU samples at the top of every iteration
    # (preStepCallback, Optimizer.cpp:801-805 -> InitCalibration.cpp:316-325): same here, on the device
    rs_note = attach_rs_imu(p, s)

    def barrier():
        if dist is not None:
            import torch
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(warmup):
        p.lm_iteration_timed(lam)
    sampler = ClockSampler(local_rank)
    sampler.start()
    # device-resident timing: K iterations, CUDA events inside the library on its launch stream
    barrier()
    phases, tot = {}, []
    wall0 = time.perf_counter()
    for _ in range(a.steps):
        t, c0, c1 = p.lm_iteration_timed(lam)
        tot.append(t["total"])
        for k_, v in t.items():
            phases[k_] = phases.get(k_, 0.0) + v / a.steps
    barrier()
    wall_ms = (time.perf_counter() - wall0) * 1e3 / a.steps
    ms_dev = float(np.mean(tot))
    # end to end through the C ABI with host buffers: variables up, iteration, variables + costs down
    host_vars = s.graph.var_data.copy()
    for _ in range(2):
        p.set_vars(host_vars); p.lm_iteration_timed(lam); _ = p.vars
    barrier()
    e0 = time.perf_counter()
    for _ in range(a.steps):
        p.set_vars(host_vars)
        p.lm_iteration_timed(lam)
        out_vars = p.vars
    barrier()
    e2e_ms = (time.perf_counter() - e0) * 1e3 / a.steps
    sampler.stop_flag = True
    sampler.join(timeout=2)
    shard = p.shard_info()
    if dist is not None:
        import torch
        names = sorted(phases)
        tt = torch.tensor([ms_dev, e2e_ms, wall_ms] + [phases[k_] for k_ in names], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)  # every timing is the max over ranks
        vals = [float(x) for x in tt.tolist()]
        ms_dev, e2e_ms, wall_ms = vals[:3]
        phases = dict(zip(names, vals[3:]))
        own = torch.tensor([shard["owned_landmarks"], shard["owned_visual"]], device="cuda", dtype=torch.float64)
        gathered = [torch.zeros_like(own) for _ in range(world)]
        dist.all_gather(gathered, own)
        shard["per_rank_landmarks"] = [int(x[0]) for x in gathered]
        shard["per_rank_observations"] = [int(x[1]) for x in gathered]
    if rank != 0:
        p.close(); ctx.close()
        if dist is not None:
            dist.destroy_process_group()
        return 0

    peaks, which = read_peaks()
    work = p.work()
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    gh_ms = phases["grad_hess"]
    dmma = ctx.dmma_peak_tflops()
    roof_gh = {"bound": "hbm", "achieved": work["visual_bytes"] / (gh_ms * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
               "traffic": None, "kernel": "gradHessVisualChunkKernel<GS>+<RS>", "peak_source": f"{which} (MEASURED_PEAKS.json hbm_gbs)",
               "algorithmic_bytes": work["visual_bytes"], "ms": gh_ms}
    roof_gh["frac"] = roof_gh["achieved"] / hbm
    fac_ms = phases["factor_schur"] + phases["factor_chol"]
    roof_fac = {"bound": "tensor", "achieved": (work["schur_flops"] + work["chol_flops"]) / (fac_ms * 1e-3) / 1e12, "peak": dmma,
                "unit": "TFLOP/s", "traffic": None, "kernel": "schurKernel + lumpFactor/lumpUpdate",
                "peak_source": "FP64 mma.sync m8n8k4 probe measured in this run (no FP64 entry in MEASURED_PEAKS.json)",
                "algorithmic_flops": work["schur_flops"] + work["chol_flops"], "ms": fac_ms}
    roof_fac["frac"] = roof_fac["achieved"] / dmma if dmma > 0 else None
    # DRAM traffic of the same kernels from the committed ncu --set full capture of this workload (N=1)
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            tr = json.load(f).get(a.workload)
        if tr and world == 1:
            roof_gh["traffic"] = tr["gradHessVisualChunkKernel<GS>+<RS>"]
            roof_gh["traffic_note"] = "bytes per iteration (one launch of each visual kernel); " + tr["source"]
            roof_fac["traffic"] = tr["schurGroupUpdateKernel"]
            roof_fac["traffic_note"] = "schurGroupUpdateKernel only, bytes per launch; " + tr["source"]
    except Exception:
        pass
    dominant = roof_gh if gh_ms >= fac_ms else roof_fac
    line = {"metric": metric, "value": ms_dev, "unit": "ms/iteration", "n_gpus": a.gpus, "steps": a.steps, "warmup": warmup,
            "ms_per_step": ms_dev, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": dict(config, observations=int(s.n_obs + s.ext.get("rs_obs", 0)), rs_observations=int(s.ext.get("rs_obs", 0)), landmarks=int(s.n_points), rigs=int(s.n_rigs), rs_refresh=rs_note,
                                                order=int(p.order), hess_doubles=int(p.data_size), lumps=int(p.num_lumps),
                                                levels=int(p.num_levels), setup_s=round(setup_s, 2)),
            "phases_ms": {k_: round(v, 4) for k_, v in phases.items()}, "wall_ms_per_step": wall_ms,
            "roofline": dominant, "roofline_other": roof_fac if dominant is roof_gh else roof_gh,
            "e2e": {"value": e2e_ms, "unit": "ms/iteration", "h2d_bytes_per_step": int(host_vars.nbytes),
                    "d2h_bytes_per_step": int(out_vars.nbytes + 8 * (64 + 4 * len(s.graph.stores)))},
            "gpu_launches": int(p.launches_per_iteration * a.steps * world), "clocks": sampler.summary()}
    if world > 1:
        per_it = coll.calls / max(1, (warmup + 2 * a.steps + 2))
        line["sharding"] = {"plan": "elimination-tree subtrees of the reduced camera system dealt to the ranks under a replicated "
                                    "shared top; landmarks (with all their observations) and factors follow the rank whose subtree "
                                    "they touch; only the shared top, the camera rhs/step and the landmark steps are summed over NCCL", "per_rank_landmarks": shard["per_rank_landmarks"],
                            "per_rank_observations": shard["per_rank_observations"],
                            "allreduce_bytes_per_iteration": int(coll.bytes / max(1, (warmup + 2 * a.steps + 2))),
                            "collectives_per_iteration": round(per_it, 2)}
    if world == 1 and not a.no_cpu_baseline:
        line["cpu_baseline"] = cpu_leg(a.workload, 2, 1)
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)
    p.close()
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
