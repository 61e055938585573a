"""Pins the oracle's Lie / camera / variable math with the reference's own test recipes
(SURVEY.md §4): exp/log round trips, analytic-vs-numeric Jacobians
(CameraProjectionTest.cpp:137-195 at point (1,1,5), tol 1e-4; TestMotionIntegral boxPlus/boxMinus)."""
import ctypes as C
import numpy as np
import pytest
from tests.oracle_lib import dptr
from visual_inertial_bundle_adjustment_b200 import synth, graph as G


def lie(lib, fn, x, n):
    x = np.ascontiguousarray(x, np.float64); out = np.zeros(n)
    assert lib.orc_lie(fn, dptr(x), dptr(out)) == 0
    return out


def test_so3_se3_roundtrip(oracle):
    rng = np.random.default_rng(42)
    for _ in range(50):
        w = rng.normal(0, 0.6, 3)
        assert np.allclose(lie(oracle, 1, lie(oracle, 0, w, 4), 3), w, atol=1e-12)
        d = rng.normal(0, 0.6, 6)
        assert np.allclose(lie(oracle, 3, lie(oracle, 2, d, 7), 6), d, atol=1e-11)
    assert np.allclose(lie(oracle, 1, lie(oracle, 0, np.zeros(3), 4), 3), 0)
    w = np.array([1e-12, -2e-12, 3e-13])
    assert np.allclose(lie(oracle, 1, lie(oracle, 0, w, 4), 3), w, rtol=1e-9)


def test_left_jacobians(oracle):
    rng = np.random.default_rng(43)
    for _ in range(20):
        w = rng.normal(0, 0.7, 3)
        J = lie(oracle, 4, w, 9).reshape(3, 3); Ji = lie(oracle, 5, w, 9).reshape(3, 3)
        assert np.allclose(J @ Ji, np.eye(3), atol=1e-12)
        # log(exp(e) exp(w)) ~ w + Jl^-1(w) e
        e = 1e-7 * rng.normal(0, 1, 3)
        q = synth.quat_mul(lie(oracle, 0, e, 4), lie(oracle, 0, w, 4))
        assert np.allclose(lie(oracle, 1, q, 3) - w, Ji @ e, atol=1e-12)
        d = rng.normal(0, 0.7, 6)
        J6 = lie(oracle, 6, d, 36).reshape(6, 6)
        e6 = 1e-7 * rng.normal(0, 1, 6)
        T = lie(oracle, 9, np.concatenate([lie(oracle, 2, e6, 7), lie(oracle, 2, d, 7)]), 7)
        assert np.allclose(lie(oracle, 3, T, 6) - d, J6 @ e6, atol=2e-12)


def test_se3_adj_inverse(oracle):
    rng = np.random.default_rng(44)
    d = rng.normal(0, 1, 6); e = 1e-6 * rng.normal(0, 1, 6)
    T = lie(oracle, 2, d, 7); Ti = lie(oracle, 8, T, 7)
    A = lie(oracle, 7, T, 36).reshape(6, 6)
    # T exp(e) T^-1 = exp(Adj(T) e)
    lhs = lie(oracle, 9, np.concatenate([lie(oracle, 9, np.concatenate([T, lie(oracle, 2, e, 7)]), 7), Ti]), 7)
    assert np.allclose(lie(oracle, 3, lhs, 6), A @ e, atol=1e-11)


def test_fisheye624_jacobians_reference_recipe(oracle):
    p = np.array([1.0, 1.0, 5.0]); prm = synth.SLAM_CAM.copy()
    uv = np.zeros(2); dP = np.zeros(6); dC = np.zeros(30)
    assert oracle.orc_fisheye624(dptr(p), dptr(prm), dptr(uv), dptr(dP), dptr(dC)) == 1
    uv_np, ok = synth.fisheye624_project(p[None], prm)
    assert ok[0] and np.allclose(uv_np[0], uv, atol=1e-12)

    def f(pp, qq):
        o = np.zeros(2); oracle.orc_fisheye624(dptr(pp), dptr(qq), dptr(o), None, None); return o
    num = np.stack([(f(p + h, prm) - f(p - h, prm)) / 2e-6 for h in 1e-6 * np.eye(3)], -1)
    assert np.allclose(num, dP.reshape(2, 3), atol=1e-4, rtol=1e-6)
    num = np.stack([(f(p, prm + h) - f(p, prm - h)) / 2e-7 for h in 1e-7 * np.eye(15)], -1)
    assert np.allclose(num, dC.reshape(2, 15), atol=1e-4, rtol=1e-6)
    z = np.array([0.1, 0.2, 1e-7])
    assert oracle.orc_fisheye624(dptr(z), dptr(prm), dptr(uv), None, None) == 0


def test_imu_boxplus_matches_generator(oracle):
    rng = np.random.default_rng(45)
    c = synth.factory_imu_calib(); s = rng.normal(0, 1e-3, 23)
    x = c.copy()
    oracle.orc_boxplus(G.VAR_IMUCALIB, synth.IMU_ALL, 0.0, dptr(x), dptr(s), 23, None)
    assert np.allclose(x, synth.imu_boxplus(c[None], s[None])[0], atol=1e-15)


def test_s2_boxplus_stays_on_sphere(oracle):
    x = np.array([0.0, 0.0, -9.81]); s = np.array([0.3, -0.2]); r = np.zeros(1)
    oracle.orc_boxplus(G.VAR_S2, 0, 9.81, dptr(x), dptr(s), 2, dptr(r))
    assert abs(np.linalg.norm(x) - 9.81) < 1e-12 and r[0] > 0
