"""bench.py contract on the CPU side: the reference arm prints exactly one JSON line with the keys
the driver reads (metric/unit/config of the b200 arm, impl, cpu_baseline, e2e with zero copies)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
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
    assert c1 < c0
    s2 = synth.make_session(**synth.CONFIGS["tiny"])
    assert bench.attach_rs_imu(Oracle(oracle, s2.graph), s2) == "no rolling-shutter camera"
