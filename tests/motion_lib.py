"""ctypes helpers over the oracle's lib/motion restatement (oracle/motion.hpp, rs_refresh.hpp, preint.hpp): the pieces the
reference's own unit tests exercise (lib/motion/preintegration/tests).  Test infrastructure only."""
import ctypes as C
import numpy as np

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int64)
_fp = C.POINTER(C.c_float)

ALL_FLAGS = 255          # ImuCalibrationOptions(true): every group incl. both time offsets -> 23 error states
NO_TIME_FLAGS = 63       # ImuCalibrationOptions::allExceptTimeOffsets() -> 21
# ImuNoiseModelParameters defaults (lib/motion/imu_types/ImuNoiseModelParameters.h:76-81)
GYRO_VAR = np.full(3, 2.7415568e-05)
ACCEL_VAR = np.full(3, 6.6297049e-3)


def d(a):
    return a.ctypes.data_as(_dp)


def factory_imu_calib():
    """factoryImuParams() (lib/motion/preintegration/ImuUtils.cpp:33-51) in the 32-double layout of the IMU calibration
    variable: gyroScale3 accelScale3 gyroBias3 accelBias3 gyroNonorth9 (column-major) accelNonorth9 (column-major) dtAccel dtGyro"""
    gN = np.array([[0.9999923706054688, -0.003904011566191912, 4.595328937284648e-05],
                   [0.003785371780395508, 0.999991774559021, -0.001455615041777492],
                   [0.0010240338742733, 0.003811037633568048, 0.9999921917915344]])
    aN = np.array([[1, 1.941762820933945e-05, -0.000217802997212857], [0, 0.9999973177909851, -0.002304504392668605], [0, 0, 1.0]])
    # normalizeImuParams (ImuUtils.cpp:14-31)
    gN[0, 0] = np.sqrt(1 - gN[0, 1] ** 2 - gN[0, 2] ** 2)
    gN[1, 1] = np.sqrt(1 - gN[1, 0] ** 2 - gN[1, 2] ** 2)
    gN[2, 2] = np.sqrt(1 - gN[2, 0] ** 2 - gN[2, 1] ** 2)
    aN[0, 0] = np.sqrt(1 - aN[0, 1] ** 2 - aN[0, 2] ** 2)
    aN[1, 1] = np.sqrt(1 - aN[1, 2] ** 2)
    c = np.zeros(32)
    c[0:3] = [0.9975922107696533, 0.9992708563804626, 1.002429008483887]
    c[3:6] = [1.00313138961792, 0.9989509582519531, 1.00210428237915]
    c[6:9] = [0.001339819049462676, 0.0001755904668243602, -0.001454736455343664]
    c[9:12] = [-0.03137952834367752, -0.1199406236410141, 0.04399538785219193]
    c[12:21] = gN.T.ravel()
    c[21:30] = aN.T.ravel()
    c[30] = 0.002755688270553946
    c[31] = 0.004112016409635544
    return c


def err_size(flags):
    return sum(n for b, n in ((1, 3), (2, 3), (4, 3), (8, 3), (16, 6), (32, 3), (64, 1), (128, 1)) if flags & b)


def idx_of(flags, bit):
    r = 0
    for b, n in ((1, 3), (2, 3), (4, 3), (8, 3), (16, 6), (32, 3), (64, 1), (128, 1)):
        if b == bit:
            return r if flags & b else -1
        if flags & b:
            r += n
    return -1


class Motion:
    def __init__(self, lib):
        self.lib = lib
        lib.orc_integrate_jac.argtypes = [_dp, _dp, C.c_double, _dp, _dp]
        lib.orc_combine.argtypes = [_dp, _dp, _dp]
        lib.orc_uncombine_left.argtypes = [_dp, _dp, _dp]
        lib.orc_combine_jacs.argtypes = [_dp, _dp, _dp, _dp, C.c_int, _dp, _dp]
        lib.orc_differentiate.argtypes = [_dp, _dp]
        lib.orc_integrate_interp.argtypes = [_dp, C.c_double, _dp]
        lib.orc_compensate_jac.argtypes = [_dp, C.c_uint32, _dp, _dp, _dp, _dp, _dp]
        lib.orc_imu_boxplus.argtypes = [_dp, _dp, C.c_uint32]
        lib.orc_preint_compute.argtypes = [C.c_int64, _ip, _fp, _fp, _dp, C.c_uint32, _dp, _dp, C.c_int64, C.c_int64, _dp, _dp, _dp, C.POINTER(C.c_int32)]
        lib.orc_integrate_measurements.argtypes = [C.c_int64, _ip, _fp, _fp, _dp, C.c_int64, C.c_int64, _dp]

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError(self.lib.orc_last_error().decode())

    def integrate(self, gyro, accel, dt, jac=False):
        g, a = np.ascontiguousarray(gyro, float), np.ascontiguousarray(accel, float)
        out, J = np.zeros(11), np.zeros(54)
        self._ck(self.lib.orc_integrate_jac(d(g), d(a), dt, d(out), d(J)))
        return (out, J.reshape(9, 6)) if jac else out

    def combine(self, a, b):
        c = np.zeros(11)
        self._ck(self.lib.orc_combine(d(a), d(b), d(c)))
        return c

    def uncombine_left(self, c, a):
        b = np.zeros(11)
        self._ck(self.lib.orc_uncombine_left(d(c), d(a), d(b)))
        return b

    def combine_jacs(self, a, b, aJ, bJ):
        n = aJ.shape[1]
        aJ, bJ = np.ascontiguousarray(aJ), np.ascontiguousarray(bJ)
        c, cJ = np.zeros(11), np.zeros((9, n))
        self._ck(self.lib.orc_combine_jacs(d(a), d(b), d(aJ), d(bJ), n, d(c), d(cJ)))
        return c, cJ

    def differentiate(self, a):
        o = np.zeros(9)
        self._ck(self.lib.orc_differentiate(d(a), d(o)))
        return o

    def integrate_interp(self, interp, dt):
        o = np.zeros(11)
        self._ck(self.lib.orc_integrate_interp(d(np.ascontiguousarray(interp)), dt, d(o)))
        return o

    def compensate(self, calib, flags, raw_gyro, raw_accel):
        E = err_size(flags)
        g, a, J = np.zeros(3), np.zeros(3), np.zeros((6, 6 + E))
        rg, ra = np.ascontiguousarray(raw_gyro, float), np.ascontiguousarray(raw_accel, float)
        self._ck(self.lib.orc_compensate_jac(d(calib), flags, d(rg), d(ra), d(g), d(a), d(J)))
        return g, a, J[:, :6].copy(), J[:, 6:].copy()

    def boxplus(self, calib, step, flags):
        c = calib.copy()
        s = np.zeros(23); s[:len(step)] = step
        self._ck(self.lib.orc_imu_boxplus(d(c), d(s), flags))
        return c

    def preint(self, time_ns, gyro, accel, calib, flags, start_us, end_us, gyro_var=GYRO_VAR, accel_var=ACCEL_VAR):
        t = np.ascontiguousarray(time_ns, np.int64)
        g = np.ascontiguousarray(gyro, np.float32); a = np.ascontiguousarray(accel, np.float32)
        rec, cov, om = np.zeros(250), np.zeros((9, 9)), np.zeros(3)
        steps = C.c_int32(0)
        gv, av = np.ascontiguousarray(gyro_var, float), np.ascontiguousarray(accel_var, float)
        self._ck(self.lib.orc_preint_compute(t.size, t.ctypes.data_as(_ip), g.ctypes.data_as(_fp), a.ctypes.data_as(_fp), d(calib), flags,
                                             d(gv), d(av), start_us, end_us, d(rec), d(cov), d(om), C.byref(steps)))
        E = err_size(flags)
        return dict(rvp=rec[:11].copy(), J=rec[11:218].reshape(9, 23)[:, :E].copy(), eval_point=rec[218:250].copy(), cov=cov, omega_at_end=om,
                    steps=steps.value, rec=rec)

    def integrate_measurements(self, time_ns, gyro, accel, calib, start_us, end_us):
        t = np.ascontiguousarray(time_ns, np.int64)
        g = np.ascontiguousarray(gyro, np.float32); a = np.ascontiguousarray(accel, np.float32)
        out = np.zeros(11)
        self._ck(self.lib.orc_integrate_measurements(t.size, t.ctypes.data_as(_ip), g.ctypes.data_as(_fp), a.ctypes.data_as(_fp), d(calib),
                                                     start_us, end_us, d(out)))
        return out


# ---- RotVelPos box operators (MotionIntegral.cpp:14-26) in numpy; quaternions are (x, y, z, w) ------------------------
def quat_mul(a, b):
    ax, ay, az, aw = a; bx, by, bz, bw = b
    return np.array([aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw, aw * bw - ax * bx - ay * by - az * bz])


def quat_conj(q):
    return np.array([-q[0], -q[1], -q[2], q[3]])


def so3_log(q):
    v, w = q[:3], q[3]
    if w < 0:
        v, w = -v, -w
    n = np.linalg.norm(v)
    if n < 1e-12:
        return 2.0 * v / w
    return 2.0 * np.arctan2(n, w) * v / n


def so3_exp(o):
    th = np.linalg.norm(o)
    if th < 1e-12:
        return np.array([0.5 * o[0], 0.5 * o[1], 0.5 * o[2], 1.0])
    s = np.sin(0.5 * th) / th
    return np.array([s * o[0], s * o[1], s * o[2], np.cos(0.5 * th)])


def box_minus(a, b):
    """boxMinus(a, b) = [log(a.R b.R^-1), a.dV - b.dV, a.dP - b.dP]"""
    return np.concatenate([so3_log(quat_mul(a[:4], quat_conj(b[:4]))), a[4:7] - b[4:7], a[7:10] - b[7:10]])


def box_plus(b, delta):
    out = b.copy()
    out[:4] = quat_mul(so3_exp(delta[:3]), b[:4])
    out[4:7] += delta[3:6]
    out[7:10] += delta[6:9]
    return out
