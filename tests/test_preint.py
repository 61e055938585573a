"""lib/motion identities of the reference's own unit tests, restated over the oracle (oracle/motion.hpp, rs_refresh.hpp,
preint.hpp) with the reference's recipes, steps and tolerances:

  TestMotionIntegral.cpp:54-175   Combine, Differentiate, Uncombine, SmallSteps (10 000-step composition), Jacobian
                                  (forward differences 1e-7, |dJ| < 1e-7), CombineJacobian (1e-10)
  TestCompensateJac.cpp:91-163    analytic vs forward-difference calibration / measurement Jacobians of the compensation
                                  (1.5e-6 / 5e-7), factoryImuParams() perturbed by boxPlus(0.1 N(0,1))
  TestPreIntegration.cpp:104-148  computePreIntegration's 9 x 23 calibration Jacobian against central differences of
                                  integrateMeasurements through boxPlus (steps 1e-6, 1e-7 for the reference-IMU time offset,
                                  3e-9 for the gyro-accel time offset; relative 1e-6 / 1e-6 / 1e-4), 1 kHz random samples,
                                  interval 0.85 s .. 1.15 s, every second case with accel and gyro time offsets aligned
  TestPreIntegration.cpp:150-203  covariance: whitened Monte-Carlo sample covariance of the integrated noise has singular
                                  values within 0.04 of 1 (fewer samples here, tolerance widened accordingly)

The reference draws from std::mt19937 + std::normal_distribution (libstdc++-defined streams, argument-evaluation order
unspecified), so the same seeds cannot reproduce its numbers: the recipes are the same, the draws are numpy's.
The device routines (csrc/dev_refresh.cuh) are checked against the same oracle in tests/test_preint_device.py."""
import numpy as np
import pytest
from tests.motion_lib import (Motion, factory_imu_calib, err_size, idx_of, box_minus, box_plus, ALL_FLAGS, NO_TIME_FLAGS,
                              GYRO_VAR, ACCEL_VAR)


@pytest.fixture(scope="module")
def mo(oracle):
    return Motion(oracle)


def rand_cap(rng, cap):
    v = rng.normal(0, 1, 3)
    n = np.linalg.norm(v)
    return v * (cap / n) if n > cap else v


def test_box_ops(mo):   # TestMotionIntegral.cpp:37-52
    rng = np.random.default_rng(42)
    for _ in range(100):
        c = mo.integrate(rng.normal(0, 1, 3), rng.normal(0, 1, 3), 1.0)
        delta = rng.normal(0, 1, 9)
        if np.linalg.norm(delta[:3]) > np.pi * 3 / 4:
            delta[:3] *= np.pi * 3 / 4 / np.linalg.norm(delta[:3])
        assert np.linalg.norm(box_minus(box_plus(c, delta), c) - delta) < 1e-10


def test_combine(mo):   # :54-67
    rng = np.random.default_rng(42)
    for _ in range(100):
        g, a, t1 = rng.normal(0, 1, 3), rng.normal(0, 1, 3), rng.uniform(0.1, 0.9)
        d = mo.combine(mo.integrate(g, a, t1), mo.integrate(g, a, 1.0 - t1))
        assert np.linalg.norm(box_minus(d, mo.integrate(g, a, 1.0))) < 1e-10


def test_differentiate(mo):   # :69-88
    rng = np.random.default_rng(42)
    for i in range(100):
        g, a = rand_cap(rng, 0.5), rng.normal(0, 1, 3)
        if i & 1:
            g *= 1e-4 / np.linalg.norm(g)
        if i & 2:
            a *= 1e-4 / np.linalg.norm(a)
        rvp = mo.integrate(g, a, 1.0)
        dpos = rand_cap(rng, 0.1)
        rvp[7:10] += dpos
        it = mo.differentiate(rvp)
        assert np.linalg.norm(g - it[:3]) < 1e-10 and np.linalg.norm(a - it[3:6]) < 1e-10 and np.linalg.norm(dpos - it[6:9]) < 1e-10
        back = mo.integrate_interp(it, 1.0)   # integrate(interp, dt) reproduces the record (MotionIntegral.cpp:111-115)
        assert np.linalg.norm(box_minus(back, rvp)) < 1e-10


def test_uncombine(mo):   # :90-110 (uncombineLeft is the one the path uses)
    rng = np.random.default_rng(42)
    for _ in range(100):
        a = mo.integrate(rand_cap(rng, 0.5), rng.normal(0, 1, 3), rng.uniform(0.1, 0.9))
        b = mo.integrate(rand_cap(rng, 0.5), rng.normal(0, 1, 3), rng.uniform(0.1, 0.9))
        rev_b = mo.uncombine_left(mo.combine(a, b), a)
        assert abs(rev_b[10] - b[10]) < 1e-10 and np.linalg.norm(box_minus(b, rev_b)) < 1e-8


def test_small_steps(mo):   # :112-127, 10 000-step composition
    rng = np.random.default_rng(42)
    for _ in range(12):
        g, a = rng.normal(0, 1, 3), rng.normal(0, 1, 3)
        steps = 10000
        step = mo.integrate(g, a, 1.0 / steps)
        acc = step
        for _ in range(1, steps):
            acc = mo.combine(acc, step)
        assert np.linalg.norm(box_minus(acc, mo.integrate(g, a, 1.0))) < 1e-10


def test_integrate_jacobian(mo):   # :129-161
    rng = np.random.default_rng(42)
    eps = 1e-7
    for _ in range(100):
        g, a, t1 = rng.normal(0, 1, 3), rng.normal(0, 1, 3), rng.uniform(0.1, 0.9)
        base, J = mo.integrate(g, a, t1, jac=True)
        num = np.zeros((9, 6))
        for i in range(6):
            gp, ap = g.copy(), a.copy()
            if i < 3:
                gp[i] += eps
            else:
                ap[i - 3] += eps
            num[:, i] = box_minus(mo.integrate(gp, ap, t1), base) / eps
        assert np.linalg.norm(num - J) < 1e-7


def test_combine_jacobian(mo):   # :163-175
    rng = np.random.default_rng(42)
    for _ in range(100):
        g, a, t1 = rng.normal(0, 1, 3), rng.normal(0, 1, 3), rng.uniform(0.1, 0.9)
        ra, aJ = mo.integrate(g, a, t1, jac=True)
        rb, bJ = mo.integrate(g, a, 1.0 - t1, jac=True)
        _, cJ = mo.integrate(g, a, 1.0, jac=True)
        _, dJ = mo.combine_jacs(ra, rb, aJ, bJ)
        assert np.linalg.norm(cJ - dJ) < 1e-10


def test_compensate_jacobian(mo):   # TestCompensateJac.cpp:91-163
    rng = np.random.default_rng(42)
    flags, E = NO_TIME_FLAGS, err_size(NO_TIME_FLAGS)
    eps = 1e-7
    worst_c = worst_m = 0.0
    for q in range(10):
        calib = factory_imu_calib()
        if q > 1:
            calib = mo.boxplus(calib, rng.normal(0, 1, E) * 0.1, flags)
        for _ in range(40):
            rg, ra = rng.normal(0, 1, 3), rng.normal(0, 1, 3)
            g0, a0, mJ, cJ = mo.compensate(calib, flags, rg, ra)
            numc = np.zeros((6, E))
            for i in range(E):
                dl = np.zeros(E); dl[i] = eps
                gp, ap, _, _ = mo.compensate(mo.boxplus(calib, dl, flags), flags, rg, ra)
                numc[:, i] = np.concatenate([gp - g0, ap - a0]) / eps
            numm = np.zeros((6, 6))
            for i in range(6):
                rgp, rap = rg.copy(), ra.copy()
                (rgp if i < 3 else rap)[i % 3] += eps
                gp, ap, _, _ = mo.compensate(calib, flags, rgp, rap)
                numm[:, i] = np.concatenate([gp - g0, ap - a0]) / eps
            worst_c = max(worst_c, np.abs(numc - cJ).max())
            worst_m = max(worst_m, np.abs(numm - mJ).max())
    assert worst_c < 1.5e-6 and worst_m < 5e-7, (worst_c, worst_m)


def gen_imu(rng, start_us, end_us):   # genImuMeasurements, TestPreIntegration.cpp:34-47: 1 kHz, accel ~ 2 g, gyro ~ pi rad/s
    t = np.arange(start_us, end_us, 1000, dtype=np.int64)
    accel = rng.normal(0, 1, (t.size, 3)) / np.sqrt(3.0) * 9.81 * 2
    gyro = rng.normal(0, 1, (t.size, 3)) / np.sqrt(3.0) * np.pi
    return t * 1000, gyro.astype(np.float32), accel.astype(np.float32)


def num_jac(mo, t, g, a, calib, flags, t0, t1):   # computeNumJac, TestPreIntegration.cpp:66-94
    E = err_size(flags)
    ref_i, ga_i = idx_of(flags, 64), idx_of(flags, 128)
    rvp0 = mo.integrate_measurements(t, g, a, calib, t0, t1)
    J = np.zeros((9, E))
    for i in range(E):
        eps = 1e-7 if i == ref_i else 3e-9 if i == ga_i else 1e-6
        dl = np.zeros(E)
        dl[i] = eps
        plus = mo.integrate_measurements(t, g, a, mo.boxplus(calib, dl, flags), t0, t1)
        minus = mo.integrate_measurements(t, g, a, mo.boxplus(calib, -dl, flags), t0, t1)
        J[:, i] = (box_minus(plus, rvp0) - box_minus(minus, rvp0)) / (2 * eps)
    return rvp0, J


def test_preintegration_jacobian(mo):   # TestPreIntegration.cpp:104-148 (250 x 5 cases there; 60 x 2 here, same recipe)
    rng = np.random.default_rng(43)
    flags, E = ALL_FLAGS, err_size(ALL_FLAGS)
    ref_i, ga_i = idx_of(flags, 64), idx_of(flags, 128)
    worst = np.zeros((9, E))
    for q in range(60):
        calib = factory_imu_calib()
        if q > 2:
            calib = mo.boxplus(calib, rng.normal(0, 1, E) * 0.05, flags)
        if q % 2 == 0:
            calib[30] = calib[31]   # accel / gyro aligned: the corner case of the time-offset Jacobian
        for _ in range(2):
            t, g, a = gen_imu(rng, 0, 2_000_000)
            res = mo.preint(t, g, a, calib, flags, 850_000, 1_150_000)
            rvp0, nj = num_jac(mo, t, g, a, calib, flags, 850_000, 1_150_000)
            assert np.linalg.norm(box_minus(res["rvp"], rvp0)) < 1e-12 and abs(res["rvp"][10] - 0.3) < 1e-12
            assert np.array_equal(res["eval_point"], calib)
            worst = np.maximum(worst, np.abs(res["J"] - nj) / np.maximum(np.abs(nj), 1.0))
    others = [i for i in range(E) if i not in (ref_i, ga_i)]
    assert worst[:, others].max() < 1e-6, worst[:, others].max()
    assert worst[:, ref_i].max() < 1e-6, worst[:, ref_i].max()
    assert worst[:, ga_i].max() < 1e-4, worst[:, ga_i].max()


def test_preintegration_covariance(mo):   # TestPreIntegration.cpp:150-203 (250 000 samples there, 6 000 here)
    flags, E = ALL_FLAGS, err_size(ALL_FLAGS)
    for q in range(3):
        rng = np.random.default_rng(39 + q)
        calib = factory_imu_calib()
        if q > 1:
            calib = mo.boxplus(calib, rng.normal(0, 1, E) * 0.005, flags)
        t, g, a = gen_imu(rng, 0, 200_000)
        res = mo.preint(t, g, a, calib, flags, 50_000, 150_000)
        Lc = np.linalg.cholesky(res["cov"])
        n, kept, acc = 6000, 0, np.zeros((9, 9))
        for _ in range(n):
            gr = (g + rng.normal(0, 1, g.shape) * np.sqrt(GYRO_VAR)).astype(np.float32)
            ar = (a + rng.normal(0, 1, a.shape) * np.sqrt(ACCEL_VAR)).astype(np.float32)
            wn = np.linalg.solve(Lc, box_minus(mo.integrate_measurements(t, gr, ar, calib, 50_000, 150_000), res["rvp"]))
            if wn @ wn > 9 + 4 * np.sqrt(2 * 9):
                continue
            acc += np.outer(wn, wn); kept += 1
        sv = np.linalg.svd(acc / kept, compute_uv=False)
        # 6 000 samples: the spread of the extreme singular values of a 9 x 9 sample covariance is ~ +-2 sqrt(9 / n) = 0.08
        assert abs(sv[0] - 1) < 0.15 and abs(sv[-1] - 1) < 0.15, sv


def test_q7_gyro_accel_offset_disabled_clobbers_raw_accel_column(mo):
    """PreIntegration.cpp:197 tests `jacInd.gyroAccelTimeOffsetIdx()` as a boolean: with the option off (index -1) the
    raw accel-z column of the step Jacobian is overwritten, which only the covariance sees.  Kept, and pinned here."""
    rng = np.random.default_rng(7)
    t, g, a = gen_imu(rng, 0, 400_000)
    calib = factory_imu_calib()
    on = mo.preint(t, g, a, calib, ALL_FLAGS, 100_000, 300_000)
    off = mo.preint(t, g, a, calib, ALL_FLAGS & ~128, 100_000, 300_000)
    assert np.allclose(on["J"][:, :22], off["J"], rtol=0, atol=1e-14)           # the calibration Jacobian is unaffected
    assert np.abs(on["cov"] - off["cov"]).max() > 1e-3 * np.abs(on["cov"]).max()  # the covariance is not
