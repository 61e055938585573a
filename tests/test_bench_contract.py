"""bench.py contract on the CPU side: the reference arm prints exactly one JSON line with the keys
the driver reads (metric/unit/config of the b200 arm, impl, cpu_baseline, e2e with zero copies)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1",
                          "--workload", "small_rs"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "ms/iteration" and d["higher_is_better"] is False
    assert d["metric"].startswith("ms per LM iteration") and "workload" in d["config"]
    assert d["value"] > 0 and d["value"] == d["cpu_baseline"]["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["sample"]
    assert d["dtype"] == "f64" and d["data"] == "synthetic" and d["vs_baseline"] is None
    # the SAME workload as the b200 arm: the whole session, every requested step, nothing scaled
    assert d["steps"] == 2 and d["warmup"] == 1 and "scale" not in d["cpu_baseline"]
    assert "small_rs" in d["config"]["workload"] and d["config"]["observations"] > 0 and d["config"]["rs_refresh"].startswith("RollingShutterData")
    runs = d["cpu_baseline"]["runs"]
    assert runs[0]["threads"] == d["cpu_baseline"]["cores"] == os.cpu_count()
    assert runs[-1]["threads"] == min(8, os.cpu_count())          # viba/common/Settings.h:87
    assert d["parity"]["cost_after"] < d["parity"]["cost_before"] and d["parity"]["step_dot_grad"] > 0
    for r in runs:
        assert abs(r["cost_before"] - d["parity"]["cost_before"]) <= 1e-9 * d["parity"]["cost_before"]


def test_default_workload_is_the_20_minute_session():
    import bench
    from visual_inertial_bundle_adjustment_b200 import synth
    assert bench.DEFAULT_WORKLOAD == "c3_20min_rs" and synth.CONFIGS[bench.DEFAULT_WORKLOAD]["duration_s"] == 1200.0


def test_session_cache_round_trip(tmp_path):
    from visual_inertial_bundle_adjustment_b200 import synth
    import numpy as np
    a, _ = synth.cached_session("tiny_rs", cache_dir=str(tmp_path))
    b, _ = synth.cached_session("tiny_rs", cache_dir=str(tmp_path))
    assert a.cache_note.startswith("generated") and b.cache_note.startswith("loaded")
    assert np.array_equal(a.graph.var_data, b.graph.var_data) and np.array_equal(a.graph.var_index, b.graph.var_index)
    for x, y in zip(a.graph.stores, b.graph.stores):
        assert x.type == y.type and np.array_equal(x.vars, y.vars) and np.array_equal(x.consts, y.consts)


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                         capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_bench_attaches_imu_samples_to_rolling_shutter_sessions(oracle):
    """the refresher is part of the timed iteration on both arms: bench.attach_rs_imu on the CPU arm's oracle"""
    import bench
    from tests.oracle_lib import Oracle
    from visual_inertial_bundle_adjustment_b200 import synth
    s = synth.make_session(**synth.CONFIGS["tiny_rs"])
    o = Oracle(oracle, s.graph)
    note = bench.attach_rs_imu(o, s)
    assert note.startswith("RollingShutterData::compute for ") and "inside every iteration" in note
    ms, c0, c1 = o.lm_iteration_timed(1e-5)   # runs the refresh first (preStepCallback)
    assert c1 < c0 and ms[7] > 0              # ms[7] = step.dot(grad) = 2 x modelCostReduction
    s2 = synth.make_session(**synth.CONFIGS["tiny"])
    assert bench.attach_rs_imu(Oracle(oracle, s2.graph), s2) == "no rolling-shutter camera"
