#!/usr/bin/env python
"""bench.py — ms per LM iteration (Jacobian+Hessian build, Schur + Cholesky solve) on synthetic
Aria-like sessions, the metric of BASELINE.json.  One "step" = one accepted-path LM iteration
(Optimizer.cpp:800-1076): RS-table refresh, grad/Hess build, damping, factor, solve, backup+apply, cost.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

Default workload: c3_20min_rs, BASELINE.json configs[2] (the 20-minute session the metric is quoted on "at
1/2/4/8 B200"); it fits one GPU.  Both arms run the SAME session: `--impl reference` times the CPU port of
the reference path (oracle/, OpenMP + OpenBLAS) on that session, no sampling, no extrapolation.

N>1 is launched by torchrun (one rank per GPU): every rank holds the same session; landmarks (with all of
their observations), factors and the subtrees of the reduced system's elimination tree are split over the
ranks; the shared top of the reduced system, the camera right-hand side / step and the landmark steps are
exchanged over NCCL at the exchange points of the C ABI's collective hook.  Total work is
fixed as N grows: "scaling": "strong".

Parity inside the line: `parity` holds cost before / cost after / step.grad of the timed iteration and their
relative differences against the CPU port (N=1, when the CPU leg runs) or against an unsharded problem rank 0
builds after the timed region (N>1); the run FAILS (exit 3) when a cost differs by more than 1e-9 relative.
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

DEFAULT_WORKLOAD = "c3_20min_rs"
METRIC = "ms per LM iteration (Jacobian+Hessian build, Schur+Cholesky solve)"
PARITY_TOL = 1e-9           # north star: per-iteration cost, FP64 relative
REFERENCE_THREADS = 8       # viba/common/Settings.h:87 (numThreads = 8)


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


METRIC_C5 = "ms per factorisation + solve of the reduced camera system (landmark Schur + supernodal Cholesky + solves; C5 sweep)"


def make_workload(name, rank=0):
    """the session of a named workload; generated once per box and cached on disk (synth.cached_session), so the
    ranks of a torchrun job and the two bench arms read the same bytes"""
    from visual_inertial_bundle_adjustment_b200 import synth
    return synth.cached_session(name, seed=42)


def headline(workload, phases):
    """the line's value from the CUDA-event phases: the whole iteration, or — on the C5 factorisation-only sweep —
    landmark Schur + supernodal Cholesky + solves (BASELINE.json configs[4])"""
    if workload.startswith("c5_"):
        return phases["factor_schur"] + phases["factor_chol"] + phases["solve"]
    return phases["total"]


def cpu_band_bytes(s, cfg):
    """LAPACK band storage the CPU stand-in would need for the reduced system: loop closures put a whole lap inside the band"""
    order = 12.0 * s.n_rigs
    lap = cfg.get("loop_period_s")
    half_bw = (12.0 * 10.0 * lap * 1.1) if lap else 2000.0
    return order * half_bw * 8.0


def workload_config(name, s, cfg, rs_note):
    """`config` of the JSON line: what the workload IS — identical on both arms and at every N"""
    rs_obs = int(s.ext.get("rs_obs", 0))
    cams = "2 global-shutter SLAM cameras" + (" + RGB rolling-shutter camera with readout/time-offset states" if cfg.get("rgb") else "")
    extra = ""
    if cfg.get("loop_period_s"):
        extra = f", closed walk with {100 * cfg.get('lc_frac', 0):g} % loop-closure tracks"
    return {"workload": f"synthetic {name}: {cfg['duration_s']:g} s Aria-like session, the same session on both arms and at every N "
                        f"({cams}, {cfg.get('n_imus', 1)} IMU(s), 5 s calibration windows, RW + factory priors{extra})",
            "observations": int(s.n_obs + rs_obs), "rs_observations": rs_obs, "landmarks": int(s.n_points), "rigs": int(s.n_rigs),
            "rs_refresh": rs_note, "lambda": 1e-5,
            "l2": "inputs larger than L2 (Hessian + factor tables exceed 126 MB); no flush needed"}


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


def attach_preint(target, session, on):
    """--recompute-preint of the reference (off by default, viba/common/Settings.h:54): raw IMU-0 samples for the IMU-0
    inertial store, so that computePreIntegration runs for every factor at the top of every iteration — on both arms"""
    if not on:
        return "off (the reference's default; --recompute-preint turns it on, on both arms)"
    from visual_inertial_bundle_adjustment_b200 import synth
    d = synth.imu0_preint_source(session)
    target.preint_set_imu(d["store"], d["time_ns"], d["gyro"], d["accel"], d["start_us"], d["end_us"], d["calib_var"], d["gyro_var"], d["accel_var"])
    target.preint_refresh(d["store"])
    return "computePreIntegration for %d IMU-0 inertial factors over %d raw samples inside every iteration (grad_hess phase)" % (d["start_us"].size, d["time_ns"].size)


def cpu_leg(s, workload, steps, warmup, thread_counts, recompute_preint=False):
    """The reference's CPU path cannot be built offline (BaSpaCho/Eigen/Sophus absent: SURVEY.md F1,F3), so
    the CPU arm is the oracle port in its baseline mode — factor loops and landmark elimination on T host
    threads (OpenMP, atomic block adds, as the reference's thread pool does with locked adds) and the
    reduced-system Cholesky through LAPACK dpbtrf on the multi-threaded OpenBLAS that ships with scipy (the
    reference: BaSpaCho's BLAS backend on OpenBLAS) — timed on the SAME session as the GPU arm, whole
    iterations, no sampling.  Returns (cpu_baseline dict for the first thread count, per-T results, rs note)."""
    from tests.oracle_lib import load_oracle, Oracle
    t0 = time.time()
    o = Oracle(load_oracle(), s.graph)
    create_s = time.time() - t0
    attach_preint(o, s, recompute_preint)
    rs_note = attach_rs_imu(o, s)
    runs = []
    for T in thread_counts:
        solver_kind = o.set_parallel(T)
        ts, c0, c1 = [], None, None
        for i in range(warmup + steps):
            ms, c0, c1 = o.lm_iteration_timed(1e-5)
            if i >= warmup:
                ts.append(ms)
        ms = np.mean(np.array(ts), 0)
        runs.append({"threads": T, "value": float(ms[2] + ms[3]) if workload.startswith("c5_") else float(ms[6]), "unit": "ms/iteration",
                     "solver": solver_kind, "steps": steps, "warmup": warmup,
                     "phases_ms": {"grad_hess": float(ms[0]), "damping": float(ms[1]), "factor": float(ms[2]), "solve": float(ms[3]),
                                   "apply": float(ms[4]), "cost": float(ms[5])},
                     "cost_before": c0, "cost_after": c1, "step_dot_grad": float(ms[7])})
    head = runs[0]
    n_obs = int(s.n_obs + s.ext.get("rs_obs", 0))
    cb = {"value": head["value"], "unit": "ms/iteration", "cores": head["threads"], "kind": "port", "solver": head["solver"],
          "sample": f"the whole {workload} session ({n_obs} observations, {s.n_rigs} rigs): {head['steps']} full iteration(s) of the oracle port after "
                    f"{head['warmup']} warm-up, {head['threads']} OpenMP threads, reduced system by {head['solver']}; not a sample, not extrapolated",
          "host_cores": os.cpu_count(), "setup_s": round(create_s, 1), "runs": runs}
    return cb, runs, rs_note


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the CPU leg of the b200 arm (and with it the N=1 parity check)")
    ap.add_argument("--recompute-preint", action="store_true", help="the reference's --recompute-preint: refresh the IMU-0 preintegration records every iteration (both arms)")
    ap.add_argument("--no-parity", action="store_true", help="N>1: skip the unsharded recomputation on rank 0")
    ap.add_argument("--reference-budget-s", type=float, default=240.0, help="reference arm: stop adding timed steps beyond this")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup

    if a.impl == "reference":
        if rank != 0:
            return 0
        s, cfg = make_workload(a.workload)
        cores = os.cpu_count() or 1
        if cpu_band_bytes(s, cfg) > 48e9:
            # loop closures >= a lap apart put the lap inside the band of the CPU stand-in's reduced-system Cholesky (LAPACK
            # dpbtrf): its storage alone exceeds host memory, and the skyline checker would need hours per factorisation
            print(json.dumps({"impl": "reference", "unavailable": "the CPU port cannot factor %s in bounded time/memory (band storage %.0f GB); "
                              "its loop-closure ingredients are timed on both arms at --workload c4_5min_lc" % (a.workload, cpu_band_bytes(s, cfg) / 1e9)}))
            return 0
        # the CPU port is deterministic work: after one warm-up iteration every iteration costs the same, so the
        # arm runs as many of the requested steps as fit the time budget (at least 1) and says how many
        from tests.oracle_lib import load_oracle, Oracle
        t0 = time.time()
        o = Oracle(load_oracle(), s.graph)
        create_s = time.time() - t0
        pre_note = attach_preint(o, s, a.recompute_preint)
        rs_note = attach_rs_imu(o, s)
        runs = []
        for T in dict.fromkeys([cores, min(REFERENCE_THREADS, cores)]):  # all host threads first (the arm's value), then the reference default
            kind = o.set_parallel(T)
            w = min(a.warmup, 1)
            t_start = time.time()
            ts, c0, c1 = [], None, None
            for i in range(w + max(1, a.steps)):
                ms, c0, c1 = o.lm_iteration_timed(1e-5)
                if i >= w:
                    ts.append(ms)
                if i >= w and time.time() - t_start > a.reference_budget_s / 2:
                    break
            m = np.mean(np.array(ts), 0)
            val = float(m[2] + m[3]) if a.workload.startswith("c5_") else float(m[6])
            runs.append({"threads": T, "value": val, "unit": "ms/iteration", "solver": kind, "steps": len(ts), "warmup": w,
                         "phases_ms": {"grad_hess": float(m[0]), "damping": float(m[1]), "factor": float(m[2]), "solve": float(m[3]),
                                       "apply": float(m[4]), "cost": float(m[5])},
                         "cost_before": c0, "cost_after": c1, "step_dot_grad": float(m[7])})
        head = runs[0]
        n_obs = int(s.n_obs + s.ext.get("rs_obs", 0))
        cb = {"value": head["value"], "unit": "ms/iteration", "cores": head["threads"], "kind": "port", "solver": head["solver"],
              "sample": f"the whole {a.workload} session ({n_obs} observations, {s.n_rigs} rigs): {head['steps']} full iteration(s) of the oracle port "
                        f"after {head['warmup']} warm-up on {head['threads']} OpenMP threads (all host threads), reduced system by {head['solver']}; "
                        f"also timed at T={runs[-1]['threads']} (the reference default, viba/common/Settings.h:87): {runs[-1]['value']:.0f} ms/iteration; "
                        "not a sample, not extrapolated",
              "host_cores": cores, "setup_s": round(create_s, 1), "runs": runs}
        line = {"impl": "reference", "metric": METRIC_C5 if a.workload.startswith("c5_") else METRIC, "value": head["value"], "unit": "ms/iteration", "n_gpus": a.gpus,
                "steps": head["steps"], "warmup": head["warmup"], "ms_per_step": head["value"], "higher_is_better": False, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": dict(workload_config(a.workload, s, cfg, rs_note), preint_refresh=pre_note),
                "cpu_baseline": cb, "parity": {"cost_before": head["cost_before"], "cost_after": head["cost_after"], "step_dot_grad": head["step_dot_grad"]},
                "e2e": {"value": head["value"], "unit": "ms/iteration", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0, "session": getattr(s, "cache_note", "")}
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
        ctx, p, coll = sharding.make_sharded_problem(s.graph, rank, world, local_rank, native=os.environ.get("VB_PY_COLLECTIVE", "0") != "1")
    else:
        ctx = capi.Context(local_rank)
        p = capi.Problem(ctx, s.graph)
    setup_s = time.time() - t0
    lam = 1e-5
    # the reference refreshes the rolling-shutter tables from the raw IMU samples at the top of every iteration
    # (preStepCallback, Optimizer.cpp:801-805 -> InitCalibration.cpp:316-325): same here, on the device
    pre_note = attach_preint(p, s, a.recompute_preint)
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
    c0 = c1 = sdg = None
    for _ in range(a.steps):
        t, c0, c1 = p.lm_iteration_timed(lam)
        sdg = t.pop("step_dot_grad")
        tot.append(headline(a.workload, t))
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
            dist.barrier()   # rank 0 may still be recomputing the unsharded iteration
            dist.destroy_process_group()
        return 0

    peaks, which = read_peaks()
    work = p.work()
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    gh_ms = phases["grad_hess"]
    dmma = ctx.dmma_peak_tflops()
    # per-GPU rooflines: the whole job's algorithmic bytes / flops over the max-rank phase time, divided by N
    roof_gh = {"bound": "hbm", "achieved": work["visual_bytes"] / (gh_ms * 1e-3) / 1e9 / world, "peak": hbm, "unit": "GB/s",
               "traffic": None, "kernel": "visual eval+assemble kernels (GS + RS)", "peak_source": f"{which} (MEASURED_PEAKS.json hbm_gbs)",
               "algorithmic_bytes": work["visual_bytes"], "ms": gh_ms, "per_gpu": True}
    roof_gh["frac"] = roof_gh["achieved"] / hbm
    fac_ms = phases["factor_schur"] + phases["factor_chol"]
    flops = work["schur_flops"] + work["chol_flops"]
    roof_fac = {"bound": "tensor", "achieved": flops / (fac_ms * 1e-3) / 1e12 / world, "peak": dmma,
                "unit": "TFLOP/s", "traffic": None, "kernel": "landmark Schur + supernodal Cholesky kernels",
                "peak_source": "builder-measured: FP64 mma.sync m8n8k4 issue-bound probe in this run (no FP64 entry in MEASURED_PEAKS.json; nominal 40)",
                "algorithmic_flops": flops, "schur_flops": work["schur_flops"], "chol_flops": work["chol_flops"], "ms": fac_ms, "per_gpu": True}
    roof_fac["frac"] = roof_fac["achieved"] / dmma if dmma > 0 else None
    low = low_fill_flops(a.workload, s)
    if low is not None:   # the flops of the low-fill (plain RCM chain) ordering: what a sequential band factorisation would need
        roof_fac["chol_flops_low_fill"] = low
        roof_fac["frac_low_fill"] = (work["schur_flops"] + low) / (fac_ms * 1e-3) / 1e12 / world / dmma if dmma > 0 else None
        roof_fac["flops_note"] = ("algorithmic_flops counts the nested-dissection ordering this library factors (parallel subtrees, more fill); "
                                  "frac_low_fill recounts the same time against the plain-RCM chain's Cholesky flops")
    # DRAM traffic of the same kernels from the committed ncu --set full capture of this workload (N=1)
    for fn in ("r02_traffic.json", "r01_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", fn)) as f:
                tr = json.load(f).get(a.workload)
            if tr and world == 1:
                roof_gh["traffic"] = tr.get("visual")
                roof_gh["traffic_note"] = "bytes per iteration (one launch of each visual kernel); " + tr["source"]
                roof_fac["traffic"] = tr.get("schur")
                roof_fac["traffic_note"] = "Schur update kernel only, bytes per launch; " + tr["source"]
                break
        except Exception:
            pass
    dominant = roof_gh if gh_ms >= fac_ms else roof_fac
    config = dict(workload_config(a.workload, s, cfg, rs_note), preint_refresh=pre_note)
    parity = {"cost_before": c0, "cost_after": c1, "step_dot_grad": sdg, "tolerance": PARITY_TOL}
    line = {"metric": METRIC_C5 if a.workload.startswith("c5_") else METRIC, "value": ms_dev, "unit": "ms/iteration", "n_gpus": a.gpus, "steps": a.steps, "warmup": warmup,
            "ms_per_step": ms_dev, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config,
            "problem": {"order": int(p.order), "hess_doubles": int(p.data_size), "lumps": int(p.num_lumps), "levels": int(p.num_levels),
                        "setup_s": round(setup_s, 2), "session": getattr(s, "cache_note", "")},
            "phases_ms": {k_: round(v, 4) for k_, v in phases.items()}, "wall_ms_per_step": wall_ms,
            "roofline": dominant, "roofline_other": roof_fac if dominant is roof_gh else roof_gh,
            "e2e": {"value": e2e_ms, "unit": "ms/iteration", "h2d_bytes_per_step": int(host_vars.nbytes),
                    "d2h_bytes_per_step": int(out_vars.nbytes + 8 * (64 + 4 * len(s.graph.stores)))},
            "gpu_launches": int(p.launches_per_iteration * a.steps * world), "clocks": sampler.summary(), "parity": parity}
    if world > 1:
        per_it = coll.calls / max(1, (warmup + 2 * a.steps + 2))
        line["sharding"] = {"plan": "elimination-tree subtrees of the reduced camera system dealt to the ranks under a shared top; landmarks "
                                    "(with all their observations) and factors follow the rank whose subtree they touch",
                            "per_rank_landmarks": shard["per_rank_landmarks"],
                            "per_rank_observations": shard["per_rank_observations"],
                            "hook": "C++ (csrc/nccl_hook.cpp: ncclAllReduce on the problem's stream)" if os.environ.get("VB_PY_COLLECTIVE", "0") != "1" else "Python (torch.distributed)",
                            "collective_bytes_per_iteration": int(coll.bytes / max(1, (warmup + 2 * a.steps + 2))),
                            "collectives_per_iteration": round(per_it, 2)}
    p.close()
    ok = True
    if world > 1 and not a.no_parity:
        # real-NCCL correctness, driver side: rank 0 recomputes the same iteration unsharded on its own GPU
        p1 = capi.Problem(ctx, s.graph)
        attach_preint(p1, s, a.recompute_preint)
        attach_rs_imu(p1, s)
        t1, u0, u1 = p1.lm_iteration_timed(lam)
        p1.close()
        parity.update({"against": "the same iteration recomputed unsharded on rank 0's GPU", "ref_cost_before": u0, "ref_cost_after": u1,
                       "ref_step_dot_grad": t1["step_dot_grad"], "rel_cost_before": rel(c0, u0), "rel_cost_after": rel(c1, u1),
                       "rel_step_dot_grad": rel(sdg, t1["step_dot_grad"])})
        ok = parity["rel_cost_before"] <= PARITY_TOL and parity["rel_cost_after"] <= PARITY_TOL and parity["rel_step_dot_grad"] <= 1e-7
    if world == 1 and not a.no_cpu_baseline and cpu_band_bytes(s, cfg) > 48e9:
        line["cpu_baseline"] = {"value": None, "unit": "ms/iteration", "cores": os.cpu_count(), "kind": "port",
                                "sample": "not run: the CPU port's band Cholesky would need %.0f GB for this loop-closure session" % (cpu_band_bytes(s, cfg) / 1e9)}
    elif world == 1 and not a.no_cpu_baseline:
        cb, runs, _ = cpu_leg(s, a.workload, 1, 0, [os.cpu_count() or 1], a.recompute_preint)
        line["cpu_baseline"] = cb
        r0 = runs[0]
        parity.update({"against": "the CPU oracle port on the same session (cpu_baseline leg)", "ref_cost_before": r0["cost_before"],
                       "ref_cost_after": r0["cost_after"], "ref_step_dot_grad": r0["step_dot_grad"],
                       "rel_cost_before": rel(c0, r0["cost_before"]), "rel_cost_after": rel(c1, r0["cost_after"]),
                       "rel_step_dot_grad": rel(sdg, r0["step_dot_grad"])})
        ok = parity["rel_cost_before"] <= PARITY_TOL and parity["rel_cost_after"] <= PARITY_TOL and parity["rel_step_dot_grad"] <= 1e-7
    parity["ok"] = bool(ok) if "against" in parity else None
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)
    ctx.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if not ok:
        print("bench.py: PARITY FAILURE " + json.dumps(parity), file=sys.stderr)
        return 3
    return 0


def low_fill_flops(workload, s):
    """Cholesky flops of the reduced camera system under the plain RCM chain (vb_symbolic_stats with nested dissection
    off): a host-only symbolic phase, cached beside the session"""
    from visual_inertial_bundle_adjustment_b200 import capi
    path = os.path.join(os.environ.get("VIBA_SESSION_CACHE", "/tmp/viba_b200_sessions"), f"{workload}_lowfill.json")
    try:
        with open(path) as f:
            return json.load(f)["chol_flops"]
    except Exception:
        pass
    try:
        st = capi.symbolic_stats(s.graph, 0, 0)
        with open(path, "w") as f:
            json.dump({"chol_flops": st["chol_flops"], "levels": st["levels"]}, f)
        return st["chol_flops"]
    except Exception:
        return None


if __name__ == "__main__":
    sys.exit(main())
