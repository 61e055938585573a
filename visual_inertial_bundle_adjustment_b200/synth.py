"""Deterministic synthetic Aria-like sessions: recordings are unavailable offline,
so benchmarks and parity tests run on sessions of the sizes BASELINE.json names.  The graph
construction rules follow viba/single_session: 5 s calibration windows
(InitCalibration.cpp:162-183), landmark-major visual factor insertion (VisualFactors.cpp:29-57),
IMU-major inertial factors (InertialFactors.cpp:72-100), RW factors chained per sensor
(RandomWalkFactors.cpp:20-152), one factory prior per calibration variable
(FactoryCalibPriors.cpp:33-145).
"""
import numpy as np
from . import graph as G

# Fisheye624 SLAM-camera vector: CameraProjectionTest.cpp:43-61
SLAM_CAM = np.array([242.0503165415735, 319.8053191211108, 240.9679064324467,
                     -0.02689165420917155, 0.1019919163316408, -0.07180862612716554,
                     0.01245898233008245, 0.001242086400994195, -0.0004149914187376608,
                     0.0005499786009926172, 0.0005819661606400815, -0.001232007804086031,
                     -0.0003868334860997226, -0.0005963449586486949, 1.071922498768078e-05])
# factoryImuParams(): lib/motion/preintegration/ImuUtils.cpp:33-51, in VarSpec<ImuCalibParam>
# data order (ImuCalibParam.cpp:214-231): gs3 as3 gb3 ab3 gN9(col-major) aN9(col-major) dtA dtG
_gN = np.array([[0.9999923706054688, -0.003904011566191912, 4.595328937284648e-05],
                [0.003785371780395508, 0.999991774559021, -0.001455615041777492],
                [0.0010240338742733, 0.003811037633568048, 0.9999921917915344]])
_aN = np.array([[1, 1.941762820933945e-05, -0.000217802997212857],
                [0, 0.9999973177909851, -0.002304504392668605], [0, 0, 1.0]])


def _normalize_nonorth(gN, aN):
    gN, aN = gN.copy(), aN.copy()
    gN[0, 0] = np.sqrt(1 - gN[0, 1] ** 2 - gN[0, 2] ** 2)
    gN[1, 1] = np.sqrt(1 - gN[1, 0] ** 2 - gN[1, 2] ** 2)
    gN[2, 2] = np.sqrt(1 - gN[2, 0] ** 2 - gN[2, 1] ** 2)
    aN[0, 0] = np.sqrt(1 - aN[0, 1] ** 2 - aN[0, 2] ** 2)
    aN[1, 1] = np.sqrt(1 - aN[1, 2] ** 2)
    aN[2, 2] = 1.0
    return gN, aN


def factory_imu_calib():
    gN, aN = _normalize_nonorth(_gN, _aN)
    return np.concatenate([
        [0.9975922107696533, 0.9992708563804626, 1.002429008483887],
        [1.00313138961792, 0.9989509582519531, 1.00210428237915],
        [0.001339819049462676, 0.0001755904668243602, -0.001454736455343664],
        [-0.03137952834367752, -0.1199406236410141, 0.04399538785219193],
        gN.T.ravel(), aN.T.ravel(), [0.002755688270553946, 0.004112016409635544]])


IMU_ALL = 255  # every ImuCalibrationOptions bit -> tangent dim 23


def imu_boxplus(c, s):
    """ImuCalibParam::boxPlus with all options on (ImuCalibParam.cpp:55-116); c (n,32), s (n,23)."""
    c = c.copy()
    c[:, 6:9] += s[:, 0:3]
    c[:, 9:12] += s[:, 3:6]
    c[:, 0:3] = 1.0 / (1.0 / c[:, 0:3] + s[:, 6:9])
    c[:, 3:6] = 1.0 / (1.0 / c[:, 3:6] + s[:, 9:12])
    cm = lambda r, cc: cc * 3 + r
    g = c[:, 12:21]
    for k, (r, cc) in enumerate([(0, 1), (0, 2), (1, 0), (1, 2), (2, 0), (2, 1)]):
        g[:, cm(r, cc)] += s[:, 12 + k]
    g[:, cm(0, 0)] = np.sqrt(1 - g[:, cm(0, 1)] ** 2 - g[:, cm(0, 2)] ** 2)
    g[:, cm(1, 1)] = np.sqrt(1 - g[:, cm(1, 0)] ** 2 - g[:, cm(1, 2)] ** 2)
    g[:, cm(2, 2)] = np.sqrt(1 - g[:, cm(2, 0)] ** 2 - g[:, cm(2, 1)] ** 2)
    a = c[:, 21:30]
    for k, (r, cc) in enumerate([(0, 1), (0, 2), (1, 2)]):
        a[:, cm(r, cc)] += s[:, 18 + k]
    a[:, cm(0, 0)] = np.sqrt(1 - a[:, cm(0, 1)] ** 2 - a[:, cm(0, 2)] ** 2)
    a[:, cm(1, 1)] = np.sqrt(1 - a[:, cm(1, 2)] ** 2)
    a[:, cm(2, 2)] = 1.0
    c[:, 31] += s[:, 21]
    c[:, 30] += s[:, 21] + s[:, 22]
    return c


# ---- quaternion helpers, memory order (x,y,z,w) -----------------------------------------------
def quat_from_rotvec(w):
    w = np.asarray(w, np.float64)
    th = np.linalg.norm(w, axis=-1, keepdims=True)
    small = th < 1e-10
    ths = np.where(small, 1.0, th)
    im = np.where(small, 0.5 - th * th / 48.0, np.sin(0.5 * ths) / ths)
    return np.concatenate([im * w, np.cos(0.5 * th)], axis=-1)


def quat_mul(a, b):
    ax, ay, az, aw = np.moveaxis(a, -1, 0)
    bx, by, bz, bw = np.moveaxis(b, -1, 0)
    return np.stack([aw * bx + ax * bw + ay * bz - az * by, aw * by + ay * bw + az * bx - ax * bz,
                     aw * bz + az * bw + ax * by - ay * bx, aw * bw - ax * bx - ay * by - az * bz], -1)


def quat_conj(q):
    return q * np.array([-1.0, -1.0, -1.0, 1.0])


def quat_rotate(q, v):
    x, y, z, w = np.moveaxis(q, -1, 0)
    u = np.stack([x, y, z], -1)
    t = 2.0 * np.cross(u, v)
    return v + w[..., None] * t + np.cross(u, t)


def se3_act(T, p):
    return quat_rotate(T[..., :4], p) + T[..., 4:]


def se3_inv(T):
    qi = quat_conj(T[..., :4])
    return np.concatenate([qi, -quat_rotate(qi, T[..., 4:])], -1)


def se3_mul(A, B):
    return np.concatenate([quat_mul(A[..., :4], B[..., :4]), quat_rotate(A[..., :4], B[..., 4:]) + A[..., 4:]], -1)


def se3_perturb(T, d):
    """exp([u, w]) * T for small d, first-order in the translation part (generator noise only)."""
    E = np.concatenate([quat_from_rotvec(d[..., 3:]), d[..., :3]], -1)
    return se3_mul(E, T)


def fisheye624_project(p, prm):
    """Forward Fisheye624 projection (FisheyeRadTanThinPrism.h:79-134), vectorised; p (n,3), prm (n,15)
    or (15,).  Returns uv (n,2) and a validity mask (z >= 1e-6: CameraModelParam.h:35-55)."""
    prm = np.broadcast_to(prm, (p.shape[0], prm.shape[-1]))
    z = p[:, 2]
    ok = z >= 1e-6
    iz = 1.0 / np.where(ok, z, 1.0)
    a, b = p[:, 0] * iz, p[:, 1] * iz
    r = np.sqrt(a * a + b * b)
    th = np.arctan(r)
    t2 = th * th
    rad = np.ones_like(th)
    ti = t2.copy()
    for i in range(6):
        rad += ti * prm[:, 3 + i]
        ti = ti * t2
    thdr = np.where(r < np.finfo(float).eps, 1.0, th / np.where(r == 0, 1.0, r))
    xr, yr = rad * thdr * a, rad * thdr * b
    rr = xr * xr + yr * yr
    p0, p1 = prm[:, 9], prm[:, 10]
    tmp = 2.0 * (xr * p0 + yr * p1)
    ud = xr + tmp * xr + rr * p0 + prm[:, 11] * rr + prm[:, 12] * rr * rr
    vd = yr + tmp * yr + rr * p1 + prm[:, 13] * rr + prm[:, 14] * rr * rr
    uv = np.stack([prm[:, 0] * ud + prm[:, 1], prm[:, 0] * vd + prm[:, 2]], -1)
    return uv, ok


# ---- ground-truth trajectory (closed form, ~1 m/s walk) -------------------------------------------
_LOOP_PERIOD = None  # set by make_session(loop_period_s=...): closed circular walk that revisits places


def _traj(t):
    """body(IMU0) pose in world: position, velocity, quaternion q_wb."""
    z = np.zeros_like(t)
    if _LOOP_PERIOD is None:
        p = np.stack([0.9 * t + 0.5 * np.sin(0.3 * t), 2.0 * np.sin(0.15 * t), 0.1 * np.sin(0.9 * t)], -1)
        v = np.stack([0.9 + 0.15 * np.cos(0.3 * t), 0.3 * np.cos(0.15 * t), 0.09 * np.cos(0.9 * t)], -1)
        q_yaw = quat_from_rotvec(np.stack([z, z, 0.3 * np.sin(0.1 * t)], -1))
    else:  # circle walked at 0.9 m/s, heading along the tangent: the same places come back every period
        om = 2 * np.pi / _LOOP_PERIOD
        R = 0.9 / om
        th = om * t
        p = np.stack([R * np.cos(th), R * np.sin(th), 0.1 * np.sin(0.9 * t)], -1)
        v = np.stack([-R * om * np.sin(th), R * om * np.cos(th), 0.09 * np.cos(0.9 * t)], -1)
        q_yaw = quat_from_rotvec(np.stack([z, z, th + np.pi / 2], -1))
    q0 = quat_from_rotvec(np.array([-np.pi / 2, 0.0, 0.0]))  # body z (optical axis) -> world +y
    q_wob = quat_from_rotvec(np.stack([0.05 * np.sin(1.1 * t), 0.05 * np.sin(0.7 * t + 1.0),
                                       0.08 * np.sin(0.9 * t + 2.0)], -1))
    q = quat_mul(quat_mul(q_yaw, np.broadcast_to(q0, q_yaw.shape)), q_wob)
    return p, v, q


def _pairs_in_window(t_l, tau, times):
    """(landmark, rig) candidate pairs with |t_rig - t_l| <= tau/2, landmark-major."""
    lo = np.searchsorted(times, t_l - 0.5 * tau, side="left")
    hi = np.searchsorted(times, t_l + 0.5 * tau, side="right")
    cnt = hi - lo
    lm = np.repeat(np.arange(t_l.size), cnt)
    start = np.repeat(lo, cnt)
    within = np.arange(cnt.sum()) - np.repeat(np.cumsum(cnt) - cnt, cnt)
    return lm, start + within


class Session:
    """What make_session returns: the Graph plus bookkeeping for tests and the bench."""
    pass


def make_session(duration_s=60.0, n_landmarks=20000, n_obs=1_000_000, seed=42, frame_hz=10.0,
                 window_s=5.0, outlier_frac=0.02, pixel_sigma=1.0 / 0.7, init_noise=1.0, rgb=False, n_imus=1,
                 rgb_obs_frac=0.35, loop_period_s=None, lc_frac=0.0):
    """Aria-Gen1-like session: 2 global-shutter SLAM cameras + IMU 0, every calibration state
    variable in 5 s windows with random-walk and factory-prior factors.
    `loop_period_s` + `lc_frac`: closed walk where that fraction of the landmarks is observed again
    on later laps (loop-closure tracks, BASELINE C4)."""
    global _LOOP_PERIOD
    _LOOP_PERIOD = loop_period_s
    try:
        return _make_session(duration_s, n_landmarks, n_obs, seed, frame_hz, window_s, outlier_frac, pixel_sigma,
                             init_noise, rgb, n_imus, rgb_obs_frac, loop_period_s, lc_frac)
    finally:
        _LOOP_PERIOD = None


def _make_session(duration_s, n_landmarks, n_obs, seed, frame_hz, window_s, outlier_frac, pixel_sigma, init_noise, rgb,
                  n_imus, rgb_obs_frac, loop_period_s, lc_frac):
    rng = np.random.default_rng(seed)
    n_rigs = int(round(duration_s * frame_hz))
    times = np.arange(n_rigs) / frame_hz
    n_cams = 2
    win = np.floor(times / window_s + 1e-9).astype(np.int64)
    n_win = int(win.max()) + 1
    p_wb, v_w, q_wb = _traj(times)
    T_wb = np.concatenate([q_wb, p_wb], -1)
    T_bw_gt = se3_inv(T_wb)

    # GT calibration per window (slow drift across windows)
    T_bc0 = [np.concatenate([quat_from_rotvec(np.array([0.0, s * 0.35, 0.0])), [s * 0.07, 0.0, 0.0]]) for s in (-1.0, 1.0)]
    T_cb_gt = np.zeros((n_cams, n_win, 7))
    cam_gt = np.zeros((n_cams, n_win, 17))
    for c in range(n_cams):
        drift = np.cumsum(rng.normal(0, 1.0, (n_win, 6)) * np.array([2e-5] * 3 + [5e-5] * 3), 0)
        T_cb_gt[c] = se3_perturb(np.broadcast_to(se3_inv(T_bc0[c]), (n_win, 7)), drift)
        cam_gt[c, :, :15] = SLAM_CAM + np.cumsum(rng.normal(0, 1.0, (n_win, 15)) * np.array([0.02, 0.02, 0.02] + [1e-5] * 12), 0)
    imu_gt = imu_boxplus(np.broadcast_to(factory_imu_calib(), (n_win, 32)),
                         np.cumsum(rng.normal(0, 1.0, (n_win, 23)) * 2e-5, 0))

    # landmarks around the path; pilot run calibrates the track length to hit n_obs
    def gen(nl, tau_mean, rg):
        t_l = rg.uniform(0, duration_s, nl)
        pl, _, ql = _traj(t_l)
        ang = rg.uniform(0, np.deg2rad(50), nl)
        az = rg.uniform(0, 2 * np.pi, nl)
        d = rg.uniform(1.5, 8.0, nl)
        dirb = np.stack([np.sin(ang) * np.cos(az), np.sin(ang) * np.sin(az), np.cos(ang)], -1)
        X = pl + quat_rotate(ql, dirb * d[:, None])
        tau = tau_mean * rg.uniform(0.5, 1.5, nl)
        lm, rig = _pairs_in_window(t_l, tau, times)
        if loop_period_s and lc_frac > 0:  # loop closures: the same landmark seen again on later laps
            again = np.nonzero(rg.random(nl) < lc_frac)[0]
            k = 1
            while k * loop_period_s < duration_s:
                lm2, rig2 = _pairs_in_window(t_l[again] + k * loop_period_s, tau[again], times)
                lm, rig = np.concatenate([lm, again[lm2]]), np.concatenate([rig, rig2])
                k += 1
            o = np.lexsort((rig, lm))
            lm, rig = lm[o], rig[o]
        lm = np.repeat(lm, n_cams)
        rig = np.repeat(rig, n_cams)
        cam = np.tile(np.arange(n_cams), lm.size // n_cams)
        pr = se3_act(T_bw_gt[rig], X[lm])
        pc = se3_act(T_cb_gt[cam, win[rig]], pr)
        uv, ok = fisheye624_project(pc, cam_gt[cam, win[rig], :15])
        ok &= (pc[:, 2] > 0.3) & (uv[:, 0] > 0) & (uv[:, 0] < 640) & (uv[:, 1] > 0) & (uv[:, 1] < 480)
        return X, lm[ok], rig[ok], cam[ok], uv[ok]

    tau0 = n_obs / (n_landmarks * n_cams * frame_hz)
    _, lm_p, _, _, _ = gen(min(2000, n_landmarks), tau0, np.random.default_rng(seed + 1))
    acc = max(lm_p.size / (min(2000, n_landmarks) * tau0 * n_cams * frame_hz), 0.05)
    X_gt, lm, rig, cam, uv = gen(n_landmarks, tau0 / acc, rng)
    # drop tracks with < 3 observations (InitPointTracks.cpp:29-65)
    cnt = np.bincount(lm, minlength=n_landmarks)
    keep_l = cnt >= 3
    new_id = np.cumsum(keep_l) - 1
    sel = keep_l[lm]
    lm, rig, cam, uv = new_id[lm[sel]], rig[sel], cam[sel], uv[sel]
    X_gt = X_gt[keep_l]
    P = X_gt.shape[0]
    n_ob = lm.size
    uv = uv + rng.normal(0, pixel_sigma, uv.shape)
    out = rng.random(n_ob) < outlier_frac
    uv[out] += rng.uniform(-30, 30, (int(out.sum()), 2))
    uv = uv.astype(np.float32).astype(np.float64)  # observations are FP32 on disk (PointObservation.h:22-23)

    # initial estimates = GT (+) noise
    s = init_noise
    drift = np.stack([0.02 * np.sin(0.05 * times + k) for k in range(3)] + [0.003 * np.sin(0.04 * times + k) for k in range(3)], -1)
    T_bw0 = se3_perturb(T_bw_gt, s * (drift + rng.normal(0, 1, (n_rigs, 6)) * np.array([2e-3] * 3 + [3e-4] * 3)))
    v0 = v_w + s * rng.normal(0, 0.01, v_w.shape)
    X0 = X_gt + s * rng.normal(0, 0.02, X_gt.shape)
    T_cb0 = se3_perturb(T_cb_gt, s * rng.normal(0, 1, (n_cams, n_win, 6)) * np.array([3e-4] * 3 + [5e-4] * 3))
    cam0 = cam_gt.copy()
    cam0[:, :, :15] += s * rng.normal(0, 1, (n_cams, n_win, 15)) * np.array([0.3, 0.3, 0.3] + [2e-4] * 6 + [5e-5] * 6)
    imu0 = imu_boxplus(imu_gt, s * rng.normal(0, 1, (n_win, 23)) * np.array([1e-4] * 3 + [2e-3] * 3 + [2e-4] * 6 + [1e-4] * 9 + [1e-4] * 2))

    g = G.Graph()
    v_pt = g.add_vars(G.VAR_VEC3, X0, 3)
    v_pose = g.add_vars(G.VAR_SE3, T_bw0, 6)
    v_vel = g.add_vars(G.VAR_VEC3, v0, 3)
    v_extr = [g.add_vars(G.VAR_SE3, T_cb0[c], 6) for c in range(n_cams)]
    v_intr = [g.add_vars(G.VAR_CAM, cam0[c], 15) for c in range(n_cams)]
    v_imu = g.add_vars(G.VAR_IMUCALIB, imu0, 23, flags=IMU_ALL)
    grav = np.array([0.0, 0.0, -9.81])
    v_g = g.add_vars(G.VAR_S2, grav, 2, aux=9.81, constant=True)  # SingleSessionAdapter.cpp:144-147
    v_extr_a, v_intr_a = np.stack(v_extr), np.stack(v_intr)

    # B1 visual factors, landmark-major then observation order (VisualFactors.cpp:29-57)
    o = np.lexsort((cam, rig, lm))
    lm, rig, cam, uv = lm[o], rig[o], cam[o], uv[o]
    vis_vars = np.stack([v_pt[lm], v_pose[rig], v_extr_a[cam, win[rig]], v_intr_a[cam, win[rig]]], -1)
    vis_c = np.concatenate([uv, np.broadcast_to(np.array([0.7, 0.0, 0.0, 0.7]), (n_ob, 4))], -1)
    g.add_store(G.FactorStore(G.F_VISUAL, vis_vars, vis_c, loss=(1.0, 3.0)))  # Constants.h:21-22

    # B4 inertial factors for IMU 0 between consecutive rigs (InertialFactors.cpp:72-100).  The
    # preintegrated measurement is synthesised from the GT trajectory so that the residual of
    # InertialFactor.cpp:23-123 is noise at GT; J(9x23) is a structured stand-in for the
    # calibration Jacobian of PreIntegration.cpp:136-275 (that refresher is a "next" row, §8f).
    prev, nxt = np.arange(n_rigs - 1), np.arange(1, n_rigs)
    nI = prev.size
    dt = times[nxt] - times[prev]
    qp, qn = T_bw_gt[prev, :4], T_bw_gt[nxt, :4]
    preR = quat_mul(qp, quat_conj(qn))
    dV = quat_rotate(qp, v_w[nxt] - v_w[prev] - grav * dt[:, None])
    dPp = T_bw_gt[prev, 4:] - quat_rotate(preR, T_bw_gt[nxt, 4:]) - quat_rotate(qp, v_w[prev] * dt[:, None] + grav * (0.5 * dt * dt)[:, None])
    sig = np.array([2e-4] * 3 + [3e-3] * 3 + [2e-4] * 3)
    noise = rng.normal(0, 1, (nI, 9)) * sig
    preR = quat_mul(quat_from_rotvec(noise[:, :3]), preR)
    dV, dPp = dV + noise[:, 3:6], dPp + noise[:, 6:9]
    PJ = np.zeros((nI, 9, 23))
    eye = np.eye(3)
    PJ[:, 0:3, 0:3] = -dt[:, None, None] * eye                  # rot / gyro bias
    PJ[:, 3:6, 3:6] = -dt[:, None, None] * eye                  # vel / accel bias
    PJ[:, 6:9, 3:6] = -(0.5 * dt * dt)[:, None, None] * eye     # pos / accel bias
    hat = lambda v: np.stack([np.stack([0 * v[:, 0], -v[:, 2], v[:, 1]], -1), np.stack([v[:, 2], 0 * v[:, 0], -v[:, 0]], -1),
                              np.stack([-v[:, 1], v[:, 0], 0 * v[:, 0]], -1)], -2)
    PJ[:, 3:6, 0:3] = 0.5 * dt[:, None, None] * hat(dV)
    PJ[:, 6:9, 0:3] = (dt / 3.0)[:, None, None] * hat(dPp)
    PJ[:, 0:3, 6:9] = 0.3 * dt[:, None, None] * eye             # gyro scale
    PJ[:, 3:6, 9:12] = 9.0 * dt[:, None, None] * eye            # accel scale
    PJ[:, 6:9, 9:12] = 4.5 * (dt * dt)[:, None, None] * eye
    PJ[:, :, 12:23] = rng.normal(0, 1, (nI, 9, 11)) * 0.2 * dt[:, None, None]
    pre = np.zeros((nI, G.PREINT_STRIDE))
    pre[:, 0:4], pre[:, 4:7], pre[:, 7:10], pre[:, 10] = preR, dV, dPp, dt
    pre[:, 11:218] = PJ.reshape(nI, -1)
    pre[:, 218:250] = imu0[win[prev]]                           # evalPoint = calibration at refresh time
    A = np.eye(9) + 0.05 * rng.normal(0, 1, (9, 9))
    cov = (A * sig) @ (A * sig).T
    Lam = np.linalg.inv(cov)
    Lam = 0.5 * (Lam + Lam.T)
    in_vars = np.stack([v_imu[win[prev]], v_pose[prev], v_vel[prev], v_pose[nxt], v_vel[nxt], np.broadcast_to(v_g, (nI,))], -1)
    g.add_store(G.FactorStore(G.F_INERTIAL, in_vars, np.broadcast_to(Lam.ravel(), (nI, 81)), loss=(np.inf, np.inf), aux=pre))

    # random walks between consecutive windows (RandomWalkFactors.cpp:20-152) and one factory prior
    # per calibration variable weighted by its rig count (FactoryCalibPriors.cpp:33-145)
    w0, w1 = np.arange(n_win - 1), np.arange(1, n_win)
    rigs_in_win = np.bincount(win, minlength=n_win).astype(np.float64)
    inflate = 100.0  # Settings.h:49-52
    if n_win > 1:
        d_imu = 1.0 / np.sqrt(np.array([1e-10] * 3 + [1e-7] * 3 + [1e-10] * 6 + [1e-11] * 9 + [1e-10] * 2) * window_s * inflate)
        g.add_store(G.FactorStore(G.F_RW_IMUCALIB, np.stack([v_imu[w0], v_imu[w1]], -1), np.broadcast_to(d_imu, (n_win - 1, 23))))
        d_cam = np.concatenate([1.0 / np.sqrt(np.array([1e-4] * 3 + [1e-10] * 12) * window_s * inflate), [0.0, 0.0]])
        g.add_store(G.FactorStore(G.F_RW_CAM, np.concatenate([np.stack([v[w0], v[w1]], -1) for v in v_intr]),
                                  np.broadcast_to(d_cam, (n_cams * (n_win - 1), 17))))
        d_ext = 1.0 / np.sqrt(np.array([1e-9] * 3 + [1e-8] * 3) * window_s * inflate)
        g.add_store(G.FactorStore(G.F_RW_SE3, np.concatenate([np.stack([v[w0], v[w1]], -1) for v in v_extr]),
                                  np.broadcast_to(d_ext, (n_cams * (n_win - 1), 6))))
    sd_imu = inflate * np.array([1e-3] * 3 + [5e-2] * 3 + [2e-3] * 6 + [1e-3] * 9 + [1e-3] * 2)
    c_pi = np.concatenate([np.sqrt(rigs_in_win)[:, None] / sd_imu, imu0], -1)
    g.add_store(G.FactorStore(G.F_PRIOR_IMUCALIB, v_imu[:, None], c_pi))
    sd_cam = inflate * np.array([1.0, 1.0, 1.0] + [1e-3] * 12)
    wgt = np.tile(np.sqrt(rigs_in_win), n_cams)[:, None]
    c_pc = np.concatenate([wgt / sd_cam, np.zeros((n_cams * n_win, 2)), cam0.reshape(-1, 17)], -1)
    g.add_store(G.FactorStore(G.F_PRIOR_CAM, np.concatenate(v_intr)[:, None], c_pc))
    sd_ext = inflate * np.array([1e-3] * 3 + [3e-3] * 3)
    c_pe = np.concatenate([wgt / sd_ext, se3_inv(T_cb0.reshape(-1, 7))], -1)
    g.add_store(G.FactorStore(G.F_PRIOR_SE3, np.concatenate(v_extr)[:, None], c_pe))

    ext = {}
    if rgb or n_imus > 1:
        ext = _extend_session(g, rng, dict(times=times, win=win, n_win=n_win, window_s=window_s, T_bw_gt=T_bw_gt, v_w=v_w,
                                           X_gt=X_gt, lm=lm, rig=rig, v_pt=v_pt, v_pose=v_pose, v_vel=v_vel, v_g=v_g,
                                           grav=grav, imu_gt=imu_gt, s=s, pixel_sigma=pixel_sigma, outlier_frac=outlier_frac,
                                           inflate=inflate, rigs_in_win=rigs_in_win), rgb, n_imus, rgb_obs_frac)
    g.finalize(v_pt)
    s_ = Session()
    s_.ext = ext
    s_.graph, s_.n_rigs, s_.n_win, s_.n_points, s_.n_obs = g, n_rigs, n_win, P, n_ob
    s_.ids = dict(point=v_pt, pose=v_pose, vel=v_vel, extr=v_extr, intr=v_intr, imu=v_imu, gravity=v_g)
    s_.gt = dict(X=X_gt, T_bw=T_bw_gt, vel=v_w, T_cb=T_cb_gt, cam=cam_gt, imu=imu_gt)
    s_.obs = dict(lm=lm, rig=rig, cam=cam, uv=uv)
    s_.times, s_.win, s_.loop_period_s = times, win, loop_period_s
    return s_


# BASELINE.json configs (C1..C3); the GS part of each
CONFIGS = {
    "tiny": dict(duration_s=4.0, n_landmarks=60, n_obs=1200, window_s=2.0),
    "small": dict(duration_s=12.0, n_landmarks=800, n_obs=24000),
    "tiny_rs": dict(duration_s=4.0, n_landmarks=60, n_obs=1200, window_s=2.0, rgb=True, n_imus=2),
    "small_rs": dict(duration_s=12.0, n_landmarks=800, n_obs=24000, rgb=True, n_imus=2),
    # loop closures (BASELINE C4 ingredient): a closed walk, 30 % of the landmarks re-observed on later laps
    "tiny_lc": dict(duration_s=6.0, n_landmarks=80, n_obs=1500, window_s=2.0, loop_period_s=3.0, lc_frac=0.3),
    "small_lc_rs": dict(duration_s=16.0, n_landmarks=800, n_obs=24000, rgb=True, n_imus=2, loop_period_s=8.0, lc_frac=0.3),
    "c1_60s": dict(duration_s=60.0, n_landmarks=20000, n_obs=1_000_000),
    "c1_60s_rs": dict(duration_s=60.0, n_landmarks=20000, n_obs=1_000_000, rgb=True, n_imus=2),
    "c2_5min_rs": dict(duration_s=300.0, n_landmarks=100000, n_obs=5_000_000, rgb=True, n_imus=2),
    "c2_5min": dict(duration_s=300.0, n_landmarks=100000, n_obs=5_000_000),
    # C4 ingredients at the C2 scale: RS camera, two IMUs and loop closures (three 100 s laps, 2 % of the
    # landmarks seen again on every later lap): 9x the Cholesky flops of c2_5min_rs through loop-closure fill
    "c4_5min_lc": dict(duration_s=300.0, n_landmarks=100000, n_obs=5_000_000, rgb=True, n_imus=2, loop_period_s=100.0, lc_frac=0.02),
    # BASELINE configs[3]: 30 min, RGB rolling-shutter camera, two IMUs with their own gyro / accelerometer reference time
    # offsets (2-4 ms, ImuUtils.cpp:46-47), a closed walk of three 10-minute laps with 2 % of the landmarks observed again
    # on every later lap (loop closures >= 600 s apart): 18 000 rigs, 450 k landmarks, ~30 M SLAM observations + RGB
    "c4_30min_gen2": dict(duration_s=1800.0, n_landmarks=450000, n_obs=30_000_000, rgb=True, n_imus=2, loop_period_s=600.0, lc_frac=0.02),
    # BASELINE configs[4], the factorisation-only sweep: reduced camera systems of 5 k - 50 k poses (walks at 10 Hz, ~500
    # observations per rig, sparse landmark field); bench.py times Schur + Cholesky + solves only on these
    "c5_5k": dict(duration_s=500.0, n_landmarks=50000, n_obs=2_500_000),
    "c5_10k": dict(duration_s=1000.0, n_landmarks=100000, n_obs=5_000_000),
    "c5_20k": dict(duration_s=2000.0, n_landmarks=200000, n_obs=10_000_000),
    "c5_50k": dict(duration_s=5000.0, n_landmarks=500000, n_obs=25_000_000),
    "c3_20min": dict(duration_s=1200.0, n_landmarks=300000, n_obs=20_000_000),
    "c3_20min_rs": dict(duration_s=1200.0, n_landmarks=300000, n_obs=20_000_000, rgb=True, n_imus=2),
}


# ---- motion integrals (lib/motion/preintegration/MotionIntegral.cpp:28-160), vectorised, used to
# build the per-rig rolling-shutter tables that RollingShutterData::compute produces on the host ----
def so3_log(q):
    n2 = np.sum(q[..., :3] ** 2, -1)
    n = np.sqrt(n2)
    w = q[..., 3]
    at = np.where(w < 0, np.arctan2(-n, -w), np.arctan2(n, w))
    f = np.where(n2 < 1e-20, 2.0 / w - (2.0 / 3.0) * n2 / (w ** 3), 2.0 * at / np.where(n == 0, 1.0, n))
    return f[..., None] * q[..., :3]


def rvp_integrate(gyro, accel, dt):
    om, up = gyro * dt[..., None], accel * dt[..., None]
    q = quat_from_rotvec(om)
    th2 = np.sum(om * om, -1); th = np.sqrt(th2); th4 = th2 * th2
    small = th < 1e-3
    ths = np.where(small, 1.0, th); th2s = np.where(small, 1.0, th2)
    m = (1.0 - np.cos(ths)) / th2s
    c1 = np.where(small, 0.5 - th2 / 24.0 + th4 / 729.0, m)  # F6 = 729 as in MotionIntegral.cpp:82
    c2 = np.where(small, 1.0 / 6.0 - th2 / 120.0 + th4 / 5040.0, (1.0 - np.sin(ths) / ths) / th2s)
    c3 = np.where(small, 1.0 / 24.0 - th2 / 729.0 + th4 / 40320.0, (0.5 - m) / th2s)
    wx = lambda v: np.cross(om, v)
    dV = up + c1[..., None] * wx(up) + c2[..., None] * wx(wx(up))
    upd = up * dt[..., None]
    dP = 0.5 * upd + c2[..., None] * wx(upd) + c3[..., None] * wx(wx(upd))
    return q, dV, dP


def rvp_differentiate(q, dV, dP, dt):
    om = so3_log(q)
    th2 = np.sum(om * om, -1); th = np.sqrt(th2)
    h = 0.5 * np.where(th < 1e-3, 1.0, th)
    q2 = np.where(th < 1e-3, 1.0 / 12.0 - th2 / (4.0 * 180.0) + th2 * th2 / (16.0 * 1890.0),
                  (1.0 - h * np.cos(h) / np.sin(h)) / np.where(th < 1e-3, 1.0, th2))
    ov = np.cross(om, dV)
    ups = dV - 0.5 * ov + q2[..., None] * np.cross(om, ov)
    gyro, accel = om / dt[..., None], ups / dt[..., None]
    _, _, rP = rvp_integrate(gyro, accel, dt)
    return np.concatenate([gyro, accel, (dP - rP) / dt[..., None]], -1)


RS_MAX_SAMPLES = 32
RS_STRIDE = 1 + RS_MAX_SAMPLES * 11 + (RS_MAX_SAMPLES - 1) * 9


def rs_tables(times, grav, half_ms=14):
    """Per-rig table of RollingShutterData (RollingShutterData.cpp:16-65): mid->t motion integrals at
    1 ms steps over +-half_ms and their interpolation data, from the closed-form trajectory."""
    tau = np.arange(-half_ms, half_ms + 1) * 1e-3
    n = tau.size
    assert n <= RS_MAX_SAMPLES
    p0, v0, q0 = _traj(times)
    t = times[:, None] + tau[None, :]
    p, v, q = _traj(t.ravel())
    p, v, q = p.reshape(-1, n, 3), v.reshape(-1, n, 3), q.reshape(-1, n, 4)
    qbw = quat_conj(q0)[:, None, :]
    qs = quat_mul(np.broadcast_to(qbw, q.shape), q)
    dV = quat_rotate(np.broadcast_to(qbw, q.shape), v - v0[:, None] - grav * tau[None, :, None])
    dP = quat_rotate(np.broadcast_to(qbw, q.shape), p - p0[:, None] - v0[:, None] * tau[None, :, None] - 0.5 * grav * (tau ** 2)[None, :, None])
    # interp[i] = differentiate(uncombineLeft(s[i+1], s[i]))
    qa, qc = qs[:, :-1], qs[:, 1:]
    bq = quat_mul(quat_conj(qa), qc)
    bdt = np.broadcast_to((tau[1:] - tau[:-1])[None, :], bq.shape[:2])
    bdV = quat_rotate(quat_conj(qa), dV[:, 1:] - dV[:, :-1])
    bdP = quat_rotate(quat_conj(qa), dP[:, 1:] - dP[:, :-1] - dV[:, :-1] * bdt[..., None])
    interp = rvp_differentiate(bq, bdV, bdP, bdt)
    tab = np.zeros((times.size, RS_STRIDE))
    tab[:, 0] = n
    smp = np.concatenate([qs, dV, dP, np.broadcast_to(tau[None, :, None], (times.size, n, 1))], -1)
    tab[:, 1:1 + n * 11] = smp.reshape(times.size, -1)
    off = 1 + RS_MAX_SAMPLES * 11
    tab[:, off:off + (n - 1) * 9] = interp.reshape(times.size, -1)
    return tab


def _imu0_raw(session, rate_hz=1000.0, margin_s=0.05, seed=7, noise=True):
    """Raw IMU-0 samples of a session (the reference reads them from the session's IMU CSV,
    lib/motion/imu_types/ImuDataFormat.h:14-23): time stamps in ns, float32 gyro (rad/s) and accelerometer (m/s^2)
    readings at `rate_hz`, each the mean of the true body rate / specific force over the sample's interval, distorted by
    the ground-truth calibration of its window (measured = scale_matrix * (true + bias), the inverse of
    getCompensatedImuMeasurement, read at the true time shifted by the window's gyro / accelerometer reference time
    offsets) plus a little white noise."""
    global _LOOP_PERIOD
    times, win = session.times, session.win
    dt = 1.0 / rate_hz
    n = int(round((times[-1] - times[0] + 2 * margin_s) * rate_hz)) + 1
    t_end = times[0] - margin_s + dt * np.arange(n)          # a sample closes the interval (t - dt, t]
    t_mid = t_end - 0.5 * dt
    h = 1e-4
    w_idx = np.clip(np.searchsorted(times, t_mid, side="right") - 1, 0, times.size - 1)
    calib = session.gt["imu"][win[w_idx]]                   # (n, 32): gs, as, gb, ab, Gn(col-major), An, dta, dtg
    # a sample stamped tau carries the gyro reading of true time tau - dtReferenceGyro and the accelerometer
    # reading of tau - dtReferenceAccel (enumIntegrationSteps subtracts them again, PreIntegration.cpp:36-43)
    t_g, t_a = t_mid - calib[:, 31], t_mid - calib[:, 30]
    _LOOP_PERIOD = session.loop_period_s
    try:
        w_body = so3_log(quat_mul(quat_conj(_traj(t_g - h)[2]), _traj(t_g + h)[2])) / (2 * h)
        a_world = (_traj(t_a + h)[1] - _traj(t_a - h)[1]) / (2 * h)
        qa = _traj(t_a)[2]
    finally:
        _LOOP_PERIOD = None
    grav = np.array([0.0, 0.0, -9.81])
    f_body = quat_rotate(quat_conj(qa), a_world - grav)
    Gn = calib[:, 12:21].reshape(n, 3, 3).transpose(0, 2, 1)
    An = calib[:, 21:30].reshape(n, 3, 3).transpose(0, 2, 1)
    gyro = calib[:, 0:3] * np.einsum("nij,nj->ni", Gn, w_body + calib[:, 6:9])
    accel = calib[:, 3:6] * np.einsum("nij,nj->ni", An, f_body + calib[:, 9:12])
    if noise:
        rng = np.random.default_rng(seed)
        gyro = gyro + rng.normal(0, np.sqrt(2.74e-5 * rate_hz) * 1e-2, gyro.shape)
        accel = accel + rng.normal(0, np.sqrt(6.63e-3 * rate_hz) * 1e-2, accel.shape)
    return np.round(t_end * 1e9).astype(np.int64), gyro.astype(np.float32), accel.astype(np.float32)


def imu0_samples(session, rate_hz=1000.0, margin_s=0.05, seed=7, noise=True):
    """Raw IMU-0 samples for the rolling-shutter refresher.  Returns dict(time_ns, gyro, accel,
    rig_mid_time_us, rig_calib_var, half_length_us) ready for vb_rs_set_imu; rigs without rolling-shutter observations
    get calib var -1."""
    g = session.graph
    times, win = session.times, session.win
    time_ns, gyro, accel = _imu0_raw(session, rate_hz, margin_s, seed, noise)
    rs_store = next(i for i, st in enumerate(g.stores) if st.type == G.F_VISUAL_RS)
    rigs = np.unique(g.stores[rs_store].consts[:, 6].astype(np.int64))
    calib_var = np.full(times.size, -1, np.int32)
    calib_var[rigs] = session.ids["imu"][win[rigs]]
    return dict(store=rs_store, time_ns=time_ns, gyro=gyro, accel=accel, rig_mid_time_us=np.round(times * 1e6).astype(np.int64),
                rig_calib_var=calib_var, half_length_us=13000)


def imu0_preint_source(session, rate_hz=1000.0, margin_s=0.05, seed=7, noise=True):
    """Raw IMU-0 samples and the integration interval of every IMU-0 inertial factor, ready for vb_preint_set_imu
    (SingleSessionAdapter::regenerateAllPreintegrationsFromImuMeasurements, viba/single_session/InertialFactors.cpp:19-70:
    interval = previous / next rig timestamps, measurement model = the calibration variable of the PREVIOUS rig's window,
    noise = ImuNoiseModelParameters defaults)."""
    g = session.graph
    times, win = session.times, session.win
    time_ns, gyro, accel = _imu0_raw(session, rate_hz, margin_s, seed, noise)
    store = next(i for i, st in enumerate(g.stores) if st.type == G.F_INERTIAL)
    n = g.stores[store].vars.shape[0]
    assert n == times.size - 1
    t_us = np.round(times * 1e6).astype(np.int64)
    return dict(store=store, time_ns=time_ns, gyro=gyro, accel=accel, start_us=t_us[:-1], end_us=t_us[1:],
                calib_var=np.ascontiguousarray(g.stores[store].vars[:, 0], np.int32),
                gyro_var=np.full(3, 2.7415568e-05), accel_var=np.full(3, 6.6297049e-3))


def _synth_preint(T_iw, vel_iw, prev, nxt, dt, grav, eval_point, rng):
    """Preintegration records (PREINT_STRIDE doubles) consistent with InertialFactor.cpp:23-123 at the
    given IMU trajectory, plus a fixed information matrix; J(9x23) is a structured stand-in."""
    nI = prev.size
    qp, qn = T_iw[prev, :4], T_iw[nxt, :4]
    preR = quat_mul(qp, quat_conj(qn))
    dV = quat_rotate(qp, vel_iw[nxt] - vel_iw[prev] - grav * dt[:, None])
    dP = T_iw[prev, 4:] - quat_rotate(preR, T_iw[nxt, 4:]) - quat_rotate(qp, vel_iw[prev] * dt[:, None] + grav * (0.5 * dt * dt)[:, None])
    sig = np.array([2e-4] * 3 + [3e-3] * 3 + [2e-4] * 3)
    noise = rng.normal(0, 1, (nI, 9)) * sig
    preR = quat_mul(quat_from_rotvec(noise[:, :3]), preR)
    dV, dP = dV + noise[:, 3:6], dP + noise[:, 6:9]
    PJ = np.zeros((nI, 9, 23))
    eye = np.eye(3)
    PJ[:, 0:3, 0:3] = -dt[:, None, None] * eye
    PJ[:, 3:6, 3:6] = -dt[:, None, None] * eye
    PJ[:, 6:9, 3:6] = -(0.5 * dt * dt)[:, None, None] * eye
    PJ[:, 0:3, 6:9] = 0.3 * dt[:, None, None] * eye
    PJ[:, 3:6, 9:12] = 9.0 * dt[:, None, None] * eye
    PJ[:, 6:9, 9:12] = 4.5 * (dt * dt)[:, None, None] * eye
    PJ[:, :, 12:23] = rng.normal(0, 1, (nI, 9, 11)) * 0.2 * dt[:, None, None]
    pre = np.zeros((nI, G.PREINT_STRIDE))
    pre[:, 0:4], pre[:, 4:7], pre[:, 7:10], pre[:, 10] = preR, dV, dP, dt
    pre[:, 11:218] = PJ.reshape(nI, -1)
    pre[:, 218:250] = eval_point
    A = np.eye(9) + 0.05 * rng.normal(0, 1, (9, 9))
    Lam = np.linalg.inv((A * sig) @ (A * sig).T)
    return pre, 0.5 * (Lam + Lam.T)


def _extend_session(g, rng, L, rgb, n_imus, rgb_obs_frac):
    """RGB rolling-shutter camera (B2 factors, 17-dim intrinsics with readout time and time offset)
    and a secondary IMU (B5/B6 inertial factors, body angular velocity variables, omega priors B15/B16,
    IMU extrinsics with RW + factory prior) — the BASELINE C2/C4 ingredients."""
    times, win, n_win, grav, s = L["times"], L["win"], L["n_win"], L["grav"], L["s"]
    T_bw_gt, v_w = L["T_bw_gt"], L["v_w"]
    n_rigs = times.size
    inflate, rigs_in_win, window_s = L["inflate"], L["rigs_in_win"], L["window_s"]
    w0, w1 = np.arange(n_win - 1), np.arange(1, n_win)
    ext = {}
    if rgb:
        H_IMG = 1408.0
        prm = np.concatenate([[610.0, 704.0, 704.0], 0.5 * SLAM_CAM[3:]])
        cam_gt = np.zeros((n_win, 17))
        cam_gt[:, :15] = prm + np.cumsum(rng.normal(0, 1.0, (n_win, 15)) * np.array([0.02] * 3 + [1e-5] * 12), 0)
        cam_gt[:, 15], cam_gt[:, 16] = 0.016, 0.003
        T_bc = np.concatenate([quat_from_rotvec(np.array([0.1, 0.0, 0.0])), [0.0, 0.02, 0.01]])
        T_cb_gt = se3_perturb(np.broadcast_to(se3_inv(T_bc), (n_win, 7)),
                              np.cumsum(rng.normal(0, 1.0, (n_win, 6)) * np.array([2e-5] * 3 + [5e-5] * 3), 0))
        pairs = np.unique(np.stack([L["lm"], L["rig"]], -1), axis=0)
        pairs = pairs[(pairs[:, 1] > 0) & (rng.random(pairs.shape[0]) < 2.0 * rgb_obs_frac)]  # XR_CHECK(rigIndex)
        lm, rig = pairs[:, 0], pairs[:, 1]
        dt = np.zeros(lm.size)
        for _ in range(4):  # rolling shutter: the row fixes the capture time, iterate to the fixed point
            p, _, q = _traj(times[rig] + dt)
            Tbw_t = se3_inv(np.concatenate([q, p], -1))
            pc = se3_act(T_cb_gt[win[rig]], se3_act(Tbw_t, L["X_gt"][lm]))
            uv, ok = fisheye624_project(pc, cam_gt[win[rig], :15])
            dt = cam_gt[win[rig], 15] * (uv[:, 1] / H_IMG - 0.5) - cam_gt[win[rig], 16]
        ok &= (pc[:, 2] > 0.3) & (uv[:, 0] > 0) & (uv[:, 0] < H_IMG) & (uv[:, 1] > 0) & (uv[:, 1] < H_IMG)
        lm, rig, uv = lm[ok], rig[ok], uv[ok]
        uv = uv + rng.normal(0, L["pixel_sigma"], uv.shape)
        outl = rng.random(lm.size) < L["outlier_frac"]
        uv[outl] += rng.uniform(-30, 30, (int(outl.sum()), 2))
        uv = uv.astype(np.float32).astype(np.float64)
        cam0 = cam_gt.copy()
        cam0[:, :15] += s * rng.normal(0, 1, (n_win, 15)) * np.array([0.3] * 3 + [2e-4] * 6 + [5e-5] * 6)
        cam0[:, 15] += s * rng.normal(0, 1e-4, n_win)
        cam0[:, 16] += s * rng.normal(0, 1e-4, n_win)
        T_cb0 = se3_perturb(T_cb_gt, s * rng.normal(0, 1, (n_win, 6)) * np.array([3e-4] * 3 + [5e-4] * 3))
        fl = G.CAM_EST_READOUT | G.CAM_EST_TIMEOFFSET | G.CAM_HAS_READOUT
        v_extr = g.add_vars(G.VAR_SE3, T_cb0, 6)
        v_intr = g.add_vars(G.VAR_CAM, cam0, 17, flags=fl)
        o = np.lexsort((rig, lm))
        lm, rig, uv = lm[o], rig[o], uv[o]
        n = lm.size
        vars_ = np.stack([L["v_pt"][lm], L["v_pose"][rig], v_extr[win[rig]], v_intr[win[rig]], L["v_vel"][rig]], -1)
        c = np.concatenate([uv, np.broadcast_to(np.array([0.7, 0.0, 0.0, 0.7]), (n, 4)), rig[:, None].astype(np.float64),
                            np.full((n, 1), H_IMG)], -1)
        aux = np.concatenate([grav, [float(n_rigs)], rs_tables(times, grav).ravel()])
        g.add_store(G.FactorStore(G.F_VISUAL_RS, vars_, c, loss=(1.0, 3.0), aux=aux))
        if n_win > 1:
            d_cam = 1.0 / np.sqrt(np.array([1e-4] * 3 + [1e-10] * 12 + [1e-12, 1e-12]) * window_s * inflate)
            g.add_store(G.FactorStore(G.F_RW_CAM, np.stack([v_intr[w0], v_intr[w1]], -1), np.broadcast_to(d_cam, (n_win - 1, 17))))
            d_ext = 1.0 / np.sqrt(np.array([1e-9] * 3 + [1e-8] * 3) * window_s * inflate)
            g.add_store(G.FactorStore(G.F_RW_SE3, np.stack([v_extr[w0], v_extr[w1]], -1), np.broadcast_to(d_ext, (n_win - 1, 6))))
        wgt = np.sqrt(rigs_in_win)[:, None]
        sd_cam = inflate * np.array([1.0] * 3 + [1e-3] * 12 + [1e-4, 1e-4])
        g.add_store(G.FactorStore(G.F_PRIOR_CAM, v_intr[:, None], np.concatenate([wgt / sd_cam, cam0], -1)))
        sd_ext = inflate * np.array([1e-3] * 3 + [3e-3] * 3)
        g.add_store(G.FactorStore(G.F_PRIOR_SE3, v_extr[:, None], np.concatenate([wgt / sd_ext, se3_inv(T_cb0)], -1)))
        ext.update(rs_obs=n, v_extr_rgb=v_extr, v_intr_rgb=v_intr, rgb_cam_gt=cam_gt)
    if n_imus > 1:
        h = 1e-4
        w_gt = so3_log(quat_mul(quat_conj(_traj(times - h)[2]), _traj(times + h)[2])) / (2 * h)  # body frame
        v_om = g.add_vars(G.VAR_VEC3, w_gt + s * rng.normal(0, 0.01, w_gt.shape), 3)
        T_bi = np.concatenate([quat_from_rotvec(np.array([0.0, 0.0, 0.02])), [-0.10, 0.0, 0.01]])
        T_ib_gt = se3_perturb(np.broadcast_to(se3_inv(T_bi), (n_win, 7)),
                              np.cumsum(rng.normal(0, 1.0, (n_win, 6)) * np.array([2e-5] * 3 + [5e-5] * 3), 0))
        T_ib0 = se3_perturb(T_ib_gt, s * rng.normal(0, 1, (n_win, 6)) * np.array([3e-4] * 3 + [5e-4] * 3))
        imu1_gt = imu_boxplus(np.broadcast_to(factory_imu_calib(), (n_win, 32)), np.cumsum(rng.normal(0, 1.0, (n_win, 23)) * 2e-5, 0))
        imu1_0 = imu_boxplus(imu1_gt, s * rng.normal(0, 1, (n_win, 23)) * np.array([1e-4] * 3 + [2e-3] * 3 + [2e-4] * 6 + [1e-4] * 9 + [1e-4] * 2))
        v_iext = g.add_vars(G.VAR_SE3, T_ib0, 6)
        v_imu1 = g.add_vars(G.VAR_IMUCALIB, imu1_0, 23, flags=IMU_ALL)
        # GT secondary-IMU trajectory: T_imu_world = T_ib * T_bw ; v = v_b + R_wb (w x t_b_i)
        T_iw = se3_mul(T_ib_gt[win], T_bw_gt)
        t_b_i = se3_inv(T_ib_gt[win])[:, 4:]
        vel_iw = v_w + quat_rotate(quat_conj(T_bw_gt[:, :4]), np.cross(w_gt, t_b_i))
        prev, nxt = np.arange(n_rigs - 1), np.arange(1, n_rigs)
        dt = times[nxt] - times[prev]
        pre, Lam = _synth_preint(T_iw, vel_iw, prev, nxt, dt, grav, imu1_0[win[prev]], rng)
        same = win[prev] == win[nxt]
        vp, vv, vg = L["v_pose"], L["v_vel"], L["v_g"]
        a, b = prev[same], nxt[same]
        g.add_store(G.FactorStore(G.F_INERTIAL_SEC,
                                  np.stack([v_imu1[win[a]], vp[a], vv[a], v_om[a], vp[b], vv[b], v_om[b], v_iext[win[a]],
                                            np.broadcast_to(vg, a.shape)], -1),
                                  np.broadcast_to(Lam.ravel(), (a.size, 81)), loss=(np.inf, np.inf), aux=pre[same]))
        a, b = prev[~same], nxt[~same]
        if a.size:
            g.add_store(G.FactorStore(G.F_INERTIAL_SEC_SPLIT,
                                      np.stack([v_imu1[win[a]], vp[a], vv[a], v_om[a], v_iext[win[a]], vp[b], vv[b], v_om[b],
                                                v_iext[win[b]], np.broadcast_to(vg, a.shape)], -1),
                                      np.broadcast_to(Lam.ravel(), (a.size, 81)), loss=(np.inf, np.inf), aux=pre[~same]))
        if n_win > 1:
            d_imu = 1.0 / np.sqrt(np.array([1e-10] * 3 + [1e-7] * 3 + [1e-10] * 6 + [1e-11] * 9 + [1e-10] * 2) * window_s * inflate)
            g.add_store(G.FactorStore(G.F_RW_IMUCALIB, np.stack([v_imu1[w0], v_imu1[w1]], -1), np.broadcast_to(d_imu, (n_win - 1, 23))))
            d_ext = 1.0 / np.sqrt(np.array([1e-9] * 3 + [1e-8] * 3) * window_s * inflate)
            g.add_store(G.FactorStore(G.F_RW_SE3, np.stack([v_iext[w0], v_iext[w1]], -1), np.broadcast_to(d_ext, (n_win - 1, 6))))
        wgt = np.sqrt(rigs_in_win)[:, None]
        sd_imu = inflate * np.array([1e-3] * 3 + [5e-2] * 3 + [2e-3] * 6 + [1e-3] * 9 + [1e-3] * 2)
        g.add_store(G.FactorStore(G.F_PRIOR_IMUCALIB, v_imu1[:, None], np.concatenate([wgt / sd_imu, imu1_0], -1)))
        sd_ext = inflate * np.array([1e-3] * 3 + [3e-3] * 3)
        g.add_store(G.FactorStore(G.F_PRIOR_SE3, v_iext[:, None], np.concatenate([wgt / sd_ext, se3_inv(T_ib0)], -1)))
        # omega priors (OmegaPriors.cpp:19-31): gyro measurements of IMU 0 and of IMU 1 in its own frame
        g.add_store(G.FactorStore(G.F_OMEGA_PRIOR, v_om[:, None], w_gt + rng.normal(0, 0.01, w_gt.shape)))
        meas1 = quat_rotate(T_ib_gt[win, :4], w_gt) + rng.normal(0, 0.01, w_gt.shape)
        g.add_store(G.FactorStore(G.F_OMEGA_PRIOR_SEC, np.stack([v_om, v_iext[win]], -1), meas1))
        ext.update(v_omega=v_om, v_imu_extr=v_iext, v_imu1=v_imu1)
    return ext


# ---- on-disk cache of generated sessions -------------------------------------------------------------
# Generating the 20-minute session takes minutes of single-threaded numpy; bench.py runs it once per box
# (every rank of a torchrun job and both bench arms share the file).  The key covers the configuration,
# the seed and this file's contents, so an edited generator never reads a stale session.
def cached_session(name, seed=42, cache_dir=None):
    import fcntl
    import hashlib
    import os
    import pickle
    import time
    cfg = dict(CONFIGS[name])
    cfg["seed"] = seed
    with open(os.path.abspath(__file__), "rb") as f:
        src = hashlib.sha256(f.read()).hexdigest()[:12]
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "graph.py"), "rb") as f:
        src += hashlib.sha256(f.read()).hexdigest()[:6]
    key = hashlib.sha256(repr(sorted(cfg.items())).encode()).hexdigest()[:12]
    cache_dir = cache_dir or os.environ.get("VIBA_SESSION_CACHE", "/tmp/viba_b200_sessions")
    os.makedirs(cache_dir, exist_ok=True)
    path = os.path.join(cache_dir, f"{name}_{key}_{src}.pkl")
    t0 = time.time()
    with open(path + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)   # one generator per box; the others wait here and then load
        try:
            if os.path.exists(path):
                fcntl.flock(lock, fcntl.LOCK_SH)
                with open(path, "rb") as f:
                    s = pickle.load(f)
                s.cache_note = f"loaded from {path} in {time.time() - t0:.1f} s"
                return s, cfg
            s = make_session(**cfg)
            s.graph._keep = []
            tmp = path + f".tmp{os.getpid()}"
            try:
                with open(tmp, "wb") as f:
                    pickle.dump(s, f, protocol=pickle.HIGHEST_PROTOCOL)
                os.replace(tmp, path)
                s.cache_note = f"generated in {time.time() - t0:.1f} s, cached at {path}"
            except OSError as ex:          # a full /tmp must not kill the bench
                s.cache_note = f"generated in {time.time() - t0:.1f} s (not cached: {ex})"
                try:
                    os.remove(tmp)
                except OSError:
                    pass
            return s, cfg
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
