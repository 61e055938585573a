"""Per-iteration refresher of the rolling-shutter tables (RollingShutterData::compute,
lib/motion/preintegration/RollingShutterData.cpp:16-65, run on the host by the reference at the top of every LM
iteration; here vb_rs_set_imu / vb_rs_refresh and the hook at the top of vb_optimize's loop).

no GPU:  the oracle's restatement (oracle/rs_refresh.hpp) against the closed-form motion of the synthetic
         trajectory, and the library's device routine (csrc/dev_refresh.cuh) compiled as host code against the
         oracle on the same samples (1e-12);
GPU:     the kernel against the oracle (tables 1e-12, cost after a refresh rel 1e-12) and a whole LM run with
         the refresh at the top of every iteration (same iteration count, final cost rel 1e-9, state 1e-6)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from tests.oracle_lib import Oracle
from visual_inertial_bundle_adjustment_b200 import synth, capi, graph as G

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STRIDE = synth.RS_STRIDE
IOFF = 1 + synth.RS_MAX_SAMPLES * 11


def session_with_imu(name, noise=True):
    s = synth.make_session(**synth.CONFIGS[name])
    return s, synth.imu0_samples(s, noise=noise)


def set_imu(obj, d):
    obj.rs_set_imu(d["store"], d["time_ns"], d["gyro"], d["accel"], d["rig_mid_time_us"], d["rig_calib_var"], d["half_length_us"])


def tables_of(aux):
    return aux[4:].reshape(-1, STRIDE)


def compare_tables(a, b, rigs, tol):
    for r in rigs:
        n = int(a[r, 0])
        assert n == int(b[r, 0]) and 15 <= n <= synth.RS_MAX_SAMPLES
        sa, sb = a[r, 1:1 + n * 11], b[r, 1:1 + n * 11]
        assert np.abs(sa - sb).max() <= tol
        ia, ib = a[r, IOFF:IOFF + (n - 1) * 9], b[r, IOFF:IOFF + (n - 1) * 9]
        assert np.abs(ia - ib).max() <= tol * max(1.0, np.abs(ib).max())


@pytest.fixture(scope="module")
def host_refresh():
    out_dir = os.path.join(ROOT, "tests", "cpp", "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "librs_refresh_host.so")
    src = os.path.join(ROOT, "tests", "cpp", "rs_refresh_host.cpp")
    hdr = os.path.join(ROOT, "visual_inertial_bundle_adjustment_b200", "csrc", "dev_refresh.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        r = subprocess.run(["g++", "-std=c++17", "-O1", "-x", "c++", "-shared", "-fPIC", "-Wall", src, "-o", so], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-3000:]
    return C.CDLL(so)


def test_oracle_refresh_follows_the_true_motion(oracle):
    """noise-free samples from the ground-truth calibration, refreshed with that calibration: the tables must
    match the closed-form mid->t motion integrals of the trajectory (synth.rs_tables) up to the 1 kHz sampling"""
    s, d = session_with_imu("tiny_rs", noise=False)
    g = s.graph
    o = Oracle(oracle, g)
    v = o.vars
    for w, vid in enumerate(s.ids["imu"]):   # put the ground-truth calibration into the variables
        v[g.var_off[vid]:g.var_off[vid] + 32] = s.gt["imu"][w]
    o.set(4, v)
    set_imu(o, d)
    o.rs_refresh()
    tab = tables_of(o.store_aux(d["store"]))
    ref = tables_of(g.stores[d["store"]].aux)
    rigs = np.flatnonzero(d["rig_calib_var"] >= 0)
    assert rigs.size > 10
    for r in rigs:
        n = int(tab[r, 0])
        smp = tab[r, 1:1 + n * 11].reshape(n, 11)
        assert np.all(np.diff(smp[:, 10]) > 0) and smp[0, 10] == pytest.approx(-0.013, abs=1e-9) and smp[-1, 10] == pytest.approx(0.013, abs=1e-9)
        nr = int(ref[r, 0])
        rs = ref[r, 1:1 + nr * 11].reshape(nr, 11)
        for row in smp:   # closed-form table is sampled every ms: compare where the times coincide
            k = np.argmin(np.abs(rs[:, 10] - row[10]))
            if abs(rs[k, 10] - row[10]) < 1e-9:
                assert np.abs(row[:4] - rs[k, :4]).max() < 1e-6      # rotation (float32 samples)
                assert np.abs(row[4:7] - rs[k, 4:7]).max() < 2e-6    # velocity increment
                assert np.abs(row[7:10] - rs[k, 7:10]).max() < 2e-8  # position increment


@pytest.mark.parametrize("name", ["tiny_rs", "small_rs"])
def test_device_routine_on_host_matches_oracle(name, oracle, host_refresh):
    s, d = session_with_imu(name)
    g = s.graph
    o = Oracle(oracle, g)
    set_imu(o, d)
    o.rs_refresh(d["store"])
    want = tables_of(o.store_aux(d["store"]))
    n_rigs = d["rig_mid_time_us"].size
    calib_off = np.where(d["rig_calib_var"] >= 0, g.var_off[np.maximum(d["rig_calib_var"], 0)], -1).astype(np.int64)
    got = np.zeros((n_rigs, STRIDE))
    ip, fp, dp = C.POINTER(C.c_int64), C.POINTER(C.c_float), C.POINTER(C.c_double)
    t, gy, ac = d["time_ns"], np.ascontiguousarray(d["gyro"]), np.ascontiguousarray(d["accel"])
    host_refresh.rf_refresh_rigs.argtypes = [ip, fp, fp, C.c_int64, dp, ip, ip, C.c_int64, C.c_int64, dp]
    st = host_refresh.rf_refresh_rigs(t.ctypes.data_as(ip), gy.ctypes.data_as(fp), ac.ctypes.data_as(fp), t.size,
                                      g.var_data.ctypes.data_as(dp), calib_off.ctypes.data_as(ip), d["rig_mid_time_us"].ctypes.data_as(ip),
                                      d["half_length_us"], n_rigs, got.ctypes.data_as(dp))
    assert st == 0
    compare_tables(got, want, np.flatnonzero(d["rig_calib_var"] >= 0), 1e-12)
    # the refreshed tables keep every rolling-shutter observation inside the sampled window
    c, stats = o.eval_cost(False)
    assert np.isfinite(c) and stats[1] == 0


def test_device_routine_irregular_samples_and_time_offsets(oracle, host_refresh):
    """BASELINE C4 ingredients: an 800 Hz IMU with jittered time stamps and separate gyro / accelerometer time
    offsets of 2-4 ms (ImuUtils.cpp:46-47), so that integration pieces end at gyro and at accel boundaries that
    do not coincide and windows start between samples; device routine (as host code) against the oracle"""
    s = synth.make_session(**synth.CONFIGS["tiny_rs"])
    d = synth.imu0_samples(s, rate_hz=800.0)
    rng = np.random.default_rng(3)
    t = d["time_ns"] + rng.integers(-100_000, 100_000, d["time_ns"].size)
    assert np.all(np.diff(t) > 0)
    d["time_ns"] = np.ascontiguousarray(t, np.int64)
    d["half_length_us"] = 12000
    g = s.graph
    o = Oracle(oracle, g)
    v = o.vars
    for w, vid in enumerate(s.ids["imu"]):
        v[g.var_off[vid] + 30] = 3.1e-3 + 1e-4 * w   # accel reference time offset
        v[g.var_off[vid] + 31] = 2.3e-3 - 1e-4 * w   # gyro reference time offset
    o.set(4, v)
    set_imu(o, d)
    o.rs_refresh()
    want = tables_of(o.store_aux(d["store"]))
    n_rigs = d["rig_mid_time_us"].size
    calib_off = np.where(d["rig_calib_var"] >= 0, g.var_off[np.maximum(d["rig_calib_var"], 0)], -1).astype(np.int64)
    got = np.zeros((n_rigs, STRIDE))
    ip, fp, dp = C.POINTER(C.c_int64), C.POINTER(C.c_float), C.POINTER(C.c_double)
    gy, ac = np.ascontiguousarray(d["gyro"]), np.ascontiguousarray(d["accel"])
    host_refresh.rf_refresh_rigs.argtypes = [ip, fp, fp, C.c_int64, dp, ip, ip, C.c_int64, C.c_int64, dp]
    st = host_refresh.rf_refresh_rigs(d["time_ns"].ctypes.data_as(ip), gy.ctypes.data_as(fp), ac.ctypes.data_as(fp), d["time_ns"].size,
                                      v.ctypes.data_as(dp), calib_off.ctypes.data_as(ip), d["rig_mid_time_us"].ctypes.data_as(ip),
                                      d["half_length_us"], n_rigs, got.ctypes.data_as(dp))
    assert st == 0
    rigs = np.flatnonzero(d["rig_calib_var"] >= 0)
    compare_tables(got, want, rigs, 1e-12)
    for r in rigs:   # the window still spans exactly +-half length around the rig's own time
        n = int(got[r, 0])
        dts = got[r, 1:1 + n * 11].reshape(n, 11)[:, 10]
        assert dts[0] == pytest.approx(-0.012, abs=1e-9) and dts[-1] == pytest.approx(0.012, abs=1e-9) and np.all(np.diff(dts) > 0)
    # too wide a window for the table: the oracle throws, the device routine reports kTooManySamples
    d2 = dict(d, half_length_us=25000)
    o2 = Oracle(oracle, g)
    set_imu(o2, d2)
    with pytest.raises(RuntimeError, match="more than 32 samples"):
        o2.rs_refresh()
    st = host_refresh.rf_refresh_rigs(d["time_ns"].ctypes.data_as(ip), gy.ctypes.data_as(fp), ac.ctypes.data_as(fp), d["time_ns"].size,
                                      v.ctypes.data_as(dp), calib_off.ctypes.data_as(ip), d["rig_mid_time_us"].ctypes.data_as(ip),
                                      25000, n_rigs, got.ctypes.data_as(dp))
    assert st == 2


def test_refresh_errors(oracle):
    s, d = session_with_imu("tiny_rs")
    o = Oracle(oracle, s.graph)
    short = dict(d)
    short["time_ns"], short["gyro"], short["accel"] = d["time_ns"][:2000], d["gyro"][:2000], d["accel"][:2000]  # samples end mid-session
    set_imu(o, short)
    with pytest.raises(RuntimeError, match="measIndex_GT"):
        o.rs_refresh()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["tiny_rs", "small_rs"])
def test_gpu_refresh_matches_oracle(name, oracle):
    s, d = session_with_imu(name)
    g = s.graph
    o = Oracle(oracle, g)
    set_imu(o, d)
    ctx = capi.Context(0)
    p = capi.Problem(ctx, g)
    set_imu(p, d)
    o.rs_refresh(); p.rs_refresh()
    rigs = np.flatnonzero(d["rig_calib_var"] >= 0)
    compare_tables(tables_of(p.store_aux(d["store"])), tables_of(o.store_aux(d["store"])), rigs, 1e-12)
    co, so = o.eval_cost(False); cg, sg = p.compute_cost(False)
    assert abs(co - cg) <= 1e-12 * abs(co) and tuple(so) == tuple(sg)
    # whole LM run, tables recomputed from the current calibration at the top of every iteration
    st = G.default_settings(6)
    summ_o, _, _ = o.optimize(st)
    summ_g = p.optimize(st)
    assert summ_g.num_iterations == summ_o.num_iterations
    assert abs(summ_g.final_cost - summ_o.final_cost) <= 1e-9 * abs(summ_o.final_cost)
    assert np.abs(p.vars - o.vars).max() <= 1e-6
    compare_tables(tables_of(p.store_aux(d["store"])), tables_of(o.store_aux(d["store"])), rigs, 1e-9)
    # samples that end mid-session: the reference throws, the ABI reports VB_ERR_NUMERIC
    short = dict(d)
    short["time_ns"], short["gyro"], short["accel"] = d["time_ns"][:2000], d["gyro"][:2000], d["accel"][:2000]
    set_imu(p, short)
    with pytest.raises(capi.VbError) as e:
        p.rs_refresh()
    assert e.value.code == 3
    p.close(); ctx.close()


def test_oracle_preintegration_of_raw_samples_matches_the_true_motion(oracle):
    """integrateMeasurements (PreIntegration.cpp:277-306, restated in oracle/rs_refresh.hpp) over consecutive rigs,
    on noise-free samples with the ground-truth calibration, against the closed form the inertial factor expects
    (InertialFactor.cpp:47-70): R = R_bw(prev) R_wb(next), dV = R_bw(prev)(v_n - v_p - g dt),
    dP = p_p - R p_n - R_bw(prev)(v_p dt + g dt^2/2) with p = translation of T_bodyImu_world."""
    s = synth.make_session(**synth.CONFIGS["tiny_rs"])
    d = synth.imu0_samples(s, noise=False)
    T, vel, grav = s.gt["T_bw"], s.gt["vel"], np.array([0.0, 0.0, -9.81])
    ip, fp, dp = C.POINTER(C.c_int64), C.POINTER(C.c_float), C.POINTER(C.c_double)
    oracle.orc_integrate_measurements.argtypes = [C.c_int64, ip, fp, fp, dp, C.c_int64, C.c_int64, dp]
    t, gy, ac, mid = d["time_ns"], np.ascontiguousarray(d["gyro"]), np.ascontiguousarray(d["accel"]), d["rig_mid_time_us"]
    for k in range(0, s.n_rigs - 1, 3):
        calib = np.ascontiguousarray(s.gt["imu"][s.win[k]])
        out = np.zeros(11)
        assert oracle.orc_integrate_measurements(t.size, t.ctypes.data_as(ip), gy.ctypes.data_as(fp), ac.ctypes.data_as(fp),
                                                 calib.ctypes.data_as(dp), int(mid[k]), int(mid[k + 1]), out.ctypes.data_as(dp)) == 0
        dt = (mid[k + 1] - mid[k]) * 1e-6
        qp, qn = T[k, :4], T[k + 1, :4]
        R = synth.quat_mul(qp, synth.quat_conj(qn))
        dV = synth.quat_rotate(qp, vel[k + 1] - vel[k] - grav * dt)
        dP = T[k, 4:] - synth.quat_rotate(R, T[k + 1, 4:]) - synth.quat_rotate(qp, vel[k] * dt + grav * (0.5 * dt * dt))
        assert out[10] == pytest.approx(dt, abs=1e-12)
        assert min(np.abs(out[:4] - R).max(), np.abs(out[:4] + R).max()) < 5e-7   # float32 samples, 1 kHz midpoint rule
        assert np.abs(out[4:7] - dV).max() < 1e-6 and np.abs(out[7:10] - dP).max() < 2e-8


def test_oracle_refresh_resnapshots_gravity(oracle):
    """updateRollingShutterData copies the CURRENT gravity variable into every RollingShutterData (InitCalibration.cpp:317,
    RollingShutterData.cpp:21): with the variable named, a changed gravity reaches the store's aux[0..2] and the cost"""
    s, d = session_with_imu("tiny_rs")
    g = s.graph
    o = Oracle(oracle, g)
    set_imu(o, d)
    gv = int(s.ids["gravity"][0])
    o.rs_refresh()
    c0 = o.eval_cost(False)[0]
    v = o.vars
    v[g.var_off[gv]:g.var_off[gv] + 3] = [0.3, -0.2, -9.80]
    o.set(4, v)
    o.rs_refresh()
    assert np.array_equal(o.store_aux(d["store"])[:3], [0.0, 0.0, -9.81])      # not named: the stored value stays
    o.rs_set_gravity_var(d["store"], gv)
    o.rs_refresh()
    assert np.array_equal(o.store_aux(d["store"])[:3], [0.3, -0.2, -9.80])
    assert o.eval_cost(False)[0] != c0


@pytest.mark.gpu
def test_gpu_refresh_resnapshots_gravity(oracle):
    s, d = session_with_imu("tiny_rs")
    g = s.graph
    gv = int(s.ids["gravity"][0])
    o = Oracle(oracle, g)
    ctx = capi.Context(0)
    p = capi.Problem(ctx, g)
    for t in (o, p):
        set_imu(t, d)
        t.rs_set_gravity_var(d["store"], gv)
    v = o.vars
    v[g.var_off[gv]:g.var_off[gv] + 3] = [0.3, -0.2, -9.80]
    o.set(4, v); p.set_vars(v)
    o.rs_refresh(); p.rs_refresh()
    assert np.array_equal(p.store_aux(d["store"])[:3], [0.3, -0.2, -9.80])
    co, cg = o.eval_cost(False)[0], p.compute_cost(False)[0]
    assert abs(co - cg) <= 1e-12 * abs(co)
    with pytest.raises(capi.VbError):
        p.rs_set_gravity_var(d["store"], int(s.ids["pose"][0]))
    p.close(); ctx.close()
