"""Session files in the reference's formats (first half), so that a synthetic session can be written where
a real `ark_vi_ba` would read it and its output read back:

  session_observations.csv   interfaces/ark/point_observation/PointObservationFormat.h:13-23 (9 columns; values are stored
                             as float32 in PointObservation and widened at factor creation, PointObservation.h:22-23)
  imu_samples_{i}.csv        lib/motion/imu_types/ImuDataFormat.h:14-23 (8 columns)
  open_loop_trajectory.csv   interfaces/ark/io/SaveDeviceTrajectory.cpp:16-37 (20 columns), the file SessionData reads
                             (interfaces/ark/session_data/SessionData.cpp:29-38) and saveOpenLoopTrajectory writes

The calibration inputs (`factory_calibration.json`, `online_calibration.jsonl`, `vrs_source_info.json`) are projectaria
MPS formats whose readers live in deps/projectaria_tools; they are not restated here.  Host I/O only:
nothing of this runs on the device or inside the iteration."""
import csv
import os

import numpy as np

from . import graph as G

OBSERVATION_COLUMNS = ["point_id", "capture_timestamp_ns", "camera_index", "projection_base_res_x", "projection_base_res_y",
                       "sqrt_h_base_res_00", "sqrt_h_base_res_01", "sqrt_h_base_res_10", "sqrt_h_base_res_11"]
IMU_COLUMNS = ["#timestamp [ns]", "temperature [degC]", "w_RS_S_x [rad s^-1]", "w_RS_S_y [rad s^-1]", "w_RS_S_z [rad s^-1]",
               "a_RS_S_x [m s^-2]", "a_RS_S_y [m s^-2]", "a_RS_S_z [m s^-2]"]
TRAJECTORY_COLUMNS = ["tracking_timestamp_us", "utc_timestamp_ns", "session_uid", "tx_odometry_device", "ty_odometry_device",
                      "tz_odometry_device", "qx_odometry_device", "qy_odometry_device", "qz_odometry_device", "qw_odometry_device",
                      "device_linear_velocity_x_odometry", "device_linear_velocity_y_odometry", "device_linear_velocity_z_odometry",
                      "angular_velocity_x_device", "angular_velocity_y_device", "angular_velocity_z_device",
                      "gravity_x_odometry", "gravity_y_odometry", "gravity_z_odometry", "quality_score"]
OBSERVATIONS_FILE = "session_observations.csv"
IMU_FILE_PATTERN = "imu_samples_{}.csv"
TRAJECTORY_FILE = "open_loop_trajectory.csv"


def session_observations(session):
    """(point_id, capture_timestamp_ns, camera_index, uv[2], sqrtH[4]) of every visual factor of a synthetic session, in
    factor order (landmark-major, VisualFactors.cpp:29-57); cameras: 0 / 1 = SLAM, 2 = RGB"""
    g = session.graph
    pose_to_rig = {int(v): i for i, v in enumerate(session.ids["pose"])}
    t_ns = np.round(session.times * 1e9).astype(np.int64)
    out = []
    for st in g.stores:
        if st.type not in (G.F_VISUAL, G.F_VISUAL_RS):
            continue
        rig = np.array([pose_to_rig[int(v)] for v in st.vars[:, 1]])
        if st.type == G.F_VISUAL:
            extr = np.asarray(session.ids["extr"])            # [camera][window] variable ids
            cam_of = {int(v): c for c in range(extr.shape[0]) for v in extr[c]}
            cam = np.array([cam_of[int(v)] for v in st.vars[:, 2]])
        else:
            cam = np.full(st.vars.shape[0], 2)
        out.append((st.vars[:, 0].astype(np.int64), t_ns[rig], cam.astype(np.int64), st.consts[:, 0:2], st.consts[:, 2:6]))
    return tuple(np.concatenate([o[k] for o in out]) for k in range(5))


def write_observations(path, point_id, t_ns, cam, uv, sqrt_h):
    with open(path, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(OBSERVATION_COLUMNS)
        uv32, h32 = np.asarray(uv, np.float32), np.asarray(sqrt_h, np.float32)   # PointObservation holds floats
        for i in range(len(point_id)):
            w.writerow([int(point_id[i]), int(t_ns[i]), int(cam[i]), repr(float(uv32[i, 0])), repr(float(uv32[i, 1])),
                        repr(float(h32[i, 0])), repr(float(h32[i, 1])), repr(float(h32[i, 2])), repr(float(h32[i, 3]))])


def read_observations(path):
    with open(path, newline="") as f:
        r = csv.reader(f)
        header = next(r)
        if header != OBSERVATION_COLUMNS:
            raise ValueError(f"{path}: unexpected columns {header}")
        rows = list(r)
    a = np.array(rows, dtype=object)
    if a.size == 0:
        z = np.zeros(0, np.int64)
        return z, z, z, np.zeros((0, 2), np.float32), np.zeros((0, 4), np.float32)
    return (a[:, 0].astype(np.int64), a[:, 1].astype(np.int64), a[:, 2].astype(np.int64),
            a[:, 3:5].astype(np.float64).astype(np.float32), a[:, 5:9].astype(np.float64).astype(np.float32))


def write_imu(path, time_ns, gyro, accel, temperature=-800.0):
    """temperature -800 = "not available" (lib/motion/preintegration/ImuTypes.h:34-42)"""
    with open(path, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(IMU_COLUMNS)
        g, a = np.asarray(gyro, np.float32), np.asarray(accel, np.float32)
        for i in range(len(time_ns)):
            w.writerow([int(time_ns[i]), temperature] + [repr(float(x)) for x in g[i]] + [repr(float(x)) for x in a[i]])


def read_imu(path):
    with open(path, newline="") as f:
        r = csv.reader(f)
        header = next(r)
        if header != IMU_COLUMNS:
            raise ValueError(f"{path}: unexpected columns {header}")
        a = np.array(list(r), dtype=np.float64)
    if a.size == 0:
        return np.zeros(0, np.int64), np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32)
    return a[:, 0].astype(np.int64), a[:, 2:5].astype(np.float32), a[:, 5:8].astype(np.float32)


def write_trajectory(path, t_us, T_world_device, vel_world, omega_device, gravity_world, session_uid="synthetic", quality=1.0):
    """T_world_device rows = quaternion (x, y, z, w) + translation, the library's SE3 layout"""
    with open(path, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(TRAJECTORY_COLUMNS)
        for i in range(len(t_us)):
            q, t = T_world_device[i, :4], T_world_device[i, 4:]
            w.writerow([int(t_us[i]), int(t_us[i]) * 1000, session_uid] + [repr(float(x)) for x in t] + [repr(float(x)) for x in q] +
                       [repr(float(x)) for x in vel_world[i]] + [repr(float(x)) for x in omega_device[i]] +
                       [repr(float(x)) for x in gravity_world] + [quality])


def read_trajectory(path):
    with open(path, newline="") as f:
        r = csv.reader(f)
        header = next(r)
        if header != TRAJECTORY_COLUMNS:
            raise ValueError(f"{path}: unexpected columns {header}")
        rows = list(r)
    t_us = np.array([int(x[0]) for x in rows], np.int64)
    num = np.array([[float(v) for v in x[3:19]] for x in rows], np.float64).reshape(len(rows), 16)
    T = np.concatenate([num[:, 3:7], num[:, 0:3]], 1)
    return dict(t_us=t_us, T_world_device=T, vel_world=num[:, 7:10], omega_device=num[:, 10:13], gravity_world=num[:, 13:16],
                session_uid=[x[2] for x in rows], quality=np.array([float(x[19]) for x in rows]))


def se3_inverse(T):
    """inverse of rows (qx, qy, qz, qw, tx, ty, tz)"""
    q = T[:, :4] * np.array([-1.0, -1.0, -1.0, 1.0])
    x, y, z, w = q.T
    t = T[:, 4:]
    R = np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
                  np.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
                  np.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], -2)
    return np.concatenate([q, -np.einsum("nij,nj->ni", R, t)], 1)


def write_session(directory, session, var_data=None):
    """the three CSV inputs / outputs of a synthetic session; `var_data` (e.g. the optimised variables) overrides the
    session's initial values for the trajectory file"""
    from . import synth
    os.makedirs(directory, exist_ok=True)
    write_observations(os.path.join(directory, OBSERVATIONS_FILE), *session_observations(session))
    time_ns, gyro, accel = synth._imu0_raw(session)
    write_imu(os.path.join(directory, IMU_FILE_PATTERN.format(0)), time_ns, gyro, accel)
    g = session.graph
    data = g.var_data if var_data is None else var_data
    pose = np.stack([data[g.var_off[v]:g.var_off[v] + 7] for v in session.ids["pose"]])      # T_bodyImu_world
    vel = np.stack([data[g.var_off[v]:g.var_off[v] + 3] for v in session.ids["vel"]])
    grav = data[g.var_off[session.ids["gravity"][0]]:g.var_off[session.ids["gravity"][0]] + 3]
    write_trajectory(os.path.join(directory, TRAJECTORY_FILE), np.round(session.times * 1e6).astype(np.int64), se3_inverse(pose), vel,
                     np.zeros_like(vel), grav)
