"""The library's device routine for computePreIntegration (csrc/dev_preint.cuh), compiled as host code, against the
oracle's restatement (oracle/preint.hpp) — no GPU needed; and, on the GPU, the kernel that runs it against the same
oracle through the C ABI (vb_preint_set_imu / vb_preint_refresh) on a synthetic session with raw IMU samples."""
import ctypes as C
import os
import subprocess
import numpy as np
import pytest
from tests.motion_lib import Motion, factory_imu_calib, err_size, ALL_FLAGS, GYRO_VAR, ACCEL_VAR
from tests.test_preint import gen_imu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_dp, _ip, _fp = C.POINTER(C.c_double), C.POINTER(C.c_int64), C.POINTER(C.c_float)


@pytest.fixture(scope="module")
def host_preint():
    out_dir = os.path.join(ROOT, "tests", "cpp", "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "libpreint_host.so")
    src = os.path.join(ROOT, "tests", "cpp", "preint_host.cpp")
    hdrs = [os.path.join(ROOT, "visual_inertial_bundle_adjustment_b200", "csrc", h) for h in ("dev_preint.cuh", "dev_refresh.cuh")]
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(p) for p in [src] + hdrs):
        r = subprocess.run(["g++", "-std=c++17", "-O1", "-x", "c++", "-shared", "-fPIC", "-Wall", src, "-o", so], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-3000:]
    lib = C.CDLL(so)
    lib.pi_preintegrate.argtypes = [_ip, _fp, _fp, C.c_int64, _dp, C.c_uint32, _dp, _dp, C.c_int64, C.c_int64, _dp, _dp, _dp]
    return lib


def device_preint(lib, t, g, a, calib, flags, t0, t1):
    rec, cov, om = np.zeros(250), np.zeros((9, 9)), np.zeros(3)
    gv, av = GYRO_VAR.copy(), ACCEL_VAR.copy()
    st = lib.pi_preintegrate(t.ctypes.data_as(_ip), g.ctypes.data_as(_fp), a.ctypes.data_as(_fp), t.size, calib.ctypes.data_as(_dp), flags,
                             gv.ctypes.data_as(_dp), av.ctypes.data_as(_dp), t0, t1, rec.ctypes.data_as(_dp), cov.ctypes.data_as(_dp), om.ctypes.data_as(_dp))
    return st, rec, cov, om


@pytest.mark.parametrize("flags", [ALL_FLAGS, ALL_FLAGS & ~128, ALL_FLAGS & ~(64 | 128), 1 | 2, 128, 1 | 2 | 64])
def test_device_routine_matches_oracle(oracle, host_preint, flags):
    mo = Motion(oracle)
    rng = np.random.default_rng(5 + flags)
    E = err_size(flags)
    for q in range(12):
        calib = factory_imu_calib()
        if q > 1:
            calib = mo.boxplus(calib, rng.normal(0, 1, E) * 0.05, flags)
        if q % 3 == 0:
            calib[30] = calib[31]        # aligned gyro / accel time offsets: both boundaries coincide
        t, g, a = gen_imu(rng, 0, 600_000)
        if q % 4 == 1:                   # jittered sample times
            t = np.sort(t + rng.integers(-200_000, 200_000, t.size)).astype(np.int64)
        t0, t1 = 150_000 + int(rng.integers(0, 999)), 400_000 + int(rng.integers(0, 999))
        ref = mo.preint(t, g, a, calib, flags, t0, t1)
        st, rec, cov, om = device_preint(host_preint, t, g, a, calib, flags, t0, t1)
        assert st == 0
        assert np.abs(rec[:11] - ref["rec"][:11]).max() <= 1e-13
        Jd, Jo = rec[11:218].reshape(9, 23), ref["rec"][11:218].reshape(9, 23)
        assert np.abs(Jd - Jo).max() <= 1e-12 * max(1.0, np.abs(Jo).max())
        assert np.array_equal(rec[218:250], calib)
        assert np.abs(cov - ref["cov"]).max() <= 1e-12 * np.abs(ref["cov"]).max()
        assert np.abs(om - ref["omega_at_end"]).max() <= 1e-14


def test_device_routine_reports_missing_margin(host_preint):
    rng = np.random.default_rng(1)
    t, g, a = gen_imu(rng, 100_000, 300_000)
    calib = factory_imu_calib()
    st, *_ = device_preint(host_preint, t, g, a, calib, ALL_FLAGS, 50_000, 200_000)     # starts before the first sample
    assert st == 1
    st, *_ = device_preint(host_preint, t, g, a, calib, ALL_FLAGS, 150_000, 400_000)    # ends after the last sample
    assert st == 1


# ---- through the C ABI on the GPU ---------------------------------------------------------------------------------------
def _attach(target, d):
    target.preint_set_imu(d["store"], d["time_ns"], d["gyro"], d["accel"], d["start_us"], d["end_us"], d["calib_var"], d["gyro_var"], d["accel_var"])


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", ["tiny_rs", "small_rs"])
def test_gpu_preint_refresh_matches_oracle(oracle, cfg):
    """vb_preint_refresh rewrites every record of the IMU-0 inertial store from the raw samples with the CURRENT
    calibration: records 1e-12 against the oracle's computePreIntegration, covariances 1e-12 relative, and the cost /
    gradient evaluated on the refreshed records agree at the usual gates"""
    from visual_inertial_bundle_adjustment_b200 import capi, synth
    from tests.oracle_lib import Oracle
    s = synth.make_session(**synth.CONFIGS[cfg])
    d = synth.imu0_preint_source(s)
    o = Oracle(oracle, s.graph)
    ctx = capi.Context(0)
    p = capi.Problem(ctx, s.graph)
    _attach(o, d); _attach(p, d)
    o.preint_refresh(); p.preint_refresh()
    n = d["start_us"].size
    ro, rg = o.store_aux(d["store"]).reshape(n, 250), p.store_aux(d["store"]).reshape(n, 250)
    assert np.abs(rg[:, :11] - ro[:, :11]).max() <= 1e-12
    assert np.abs(rg[:, 11:218] - ro[:, 11:218]).max() <= 1e-12 * max(1.0, np.abs(ro[:, 11:218]).max())
    assert np.array_equal(rg[:, 218:], ro[:, 218:])
    co, oo = o.preint_download(d["store"], n)
    cg, og = p.preint_download(d["store"], n)
    assert np.abs(cg - co).max() <= 1e-12 * np.abs(co).max() and np.abs(og - oo).max() <= 1e-13
    c_o, c_g = o.eval_grad_hess(), p.compute_grad_hess()
    assert abs(c_g - c_o) <= 1e-12 * abs(c_o)
    from tests.test_gpu_parity import to_reg
    go, gg = o.grad, to_reg(p, o, p.grad)
    assert np.abs(gg - go).max() <= 1e-10 * np.abs(go).max()
    p.close(); ctx.close()


@pytest.mark.gpu
def test_gpu_lm_with_preint_and_rs_refresh_in_the_loop(oracle):
    """the reference's preStepCallback under --recompute-preint (main_AriaKit_ViBa.cpp:95-101): preintegrations, then the
    rolling-shutter tables, at the top of every iteration — same trace as the oracle driven the same way"""
    from visual_inertial_bundle_adjustment_b200 import capi, synth, graph as G
    from tests.oracle_lib import Oracle
    from tests.test_lm_branches import assert_traces_match
    s = synth.make_session(**synth.CONFIGS["tiny_rs"])
    d, r = synth.imu0_preint_source(s), synth.imu0_samples(s)
    st = G.default_settings(50); st.max_num_iterations = 6
    o = Oracle(oracle, s.graph)
    _attach(o, d)
    o.rs_set_imu(r["store"], r["time_ns"], r["gyro"], r["accel"], r["rig_mid_time_us"], r["rig_calib_var"], r["half_length_us"])
    so, to = o.optimize_ex(st)
    ctx = capi.Context(0)
    p = capi.Problem(ctx, s.graph)
    _attach(p, d)
    p.rs_set_imu(r["store"], r["time_ns"], r["gyro"], r["accel"], r["rig_mid_time_us"], r["rig_calib_var"], r["half_length_us"])
    sg, tg = p.optimize_ex(st)
    assert_traces_match(tg, to)
    assert sg.num_iterations == so.num_iterations and abs(sg.final_cost - so.final_cost) <= 1e-9 * abs(so.final_cost)
    assert np.abs(p.vars - o.vars).max() <= 1e-6
    # the records the last iteration used were computed from the calibration of that iteration, not the initial one
    rec = p.store_aux(d["store"]).reshape(-1, 250)
    assert not np.array_equal(rec[:, 218:], s.graph.stores[d["store"]].aux.reshape(-1, 250)[:, 218:])
    p.preint_clear_imu(d["store"])
    with pytest.raises(Exception):
        p.preint_refresh(d["store"])
    p.close(); ctx.close()


@pytest.mark.gpu
def test_gpu_preint_interval_off_the_samples_is_an_error(oracle):
    from visual_inertial_bundle_adjustment_b200 import capi, synth
    s = synth.make_session(**synth.CONFIGS["tiny_rs"])
    d = synth.imu0_preint_source(s)
    ctx = capi.Context(0)
    p = capi.Problem(ctx, s.graph)
    k = d["time_ns"].size // 2
    p.preint_set_imu(d["store"], d["time_ns"][:k], d["gyro"][:k], d["accel"][:k], d["start_us"], d["end_us"], d["calib_var"], d["gyro_var"], d["accel_var"])
    with pytest.raises(Exception, match="not enough IMU samples"):
        p.preint_refresh()
    p.close(); ctx.close()
