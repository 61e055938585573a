"""Session files in the reference's formats (first half), so that a synthetic session can be written where
a real `ark_vi_ba` would read it and its output read back:

  session_observations.csv   interfaces/ark/point_observation/PointObservationFormat.h:13-23 (9 columns; values are stored
                             as float32 in PointObservation and widened at factor creation, PointObservation.h:22-23)
  imu_samples_{label}.csv    lib/motion/imu_types/ImuDataFormat.h:14-23 (8 columns); named after the SLAM IMU label
                             (SessionData.cpp:33,323-331)
  open_loop_trajectory.csv   interfaces/ark/io/SaveDeviceTrajectory.cpp:16-37 (20 columns), the file SessionData reads
                             (interfaces/ark/session_data/SessionData.cpp:29-38) and saveOpenLoopTrajectory writes

The calibration inputs (`factory_calibration.json`, `online_calibration.jsonl`, `vrs_source_info.json`, the optional
`reset_events.json`) are projectaria MPS formats; their schemas are restated in the second half of this file from the
readers in deps/projectaria_tools, together with the calibration half of SessionData::load (`load_session_calibration`).
Host I/O only: nothing of this runs on the device or inside the iteration."""
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
    write_imu(os.path.join(directory, IMU_FILE_PATTERN.format(IMU_LABELS[0])), time_ns, gyro, accel)
    write_session_calibration(directory, session, var_data, imu_ids=[IMU_LABELS[0]])   # raw samples exist for the body IMU only
    g = session.graph
    data = g.var_data if var_data is None else var_data
    pose = np.stack([data[g.var_off[v]:g.var_off[v] + 7] for v in session.ids["pose"]])      # T_bodyImu_world
    vel = np.stack([data[g.var_off[v]:g.var_off[v] + 3] for v in session.ids["vel"]])
    grav = data[g.var_off[session.ids["gravity"][0]]:g.var_off[session.ids["gravity"][0]] + 3]
    write_trajectory(os.path.join(directory, TRAJECTORY_FILE), np.round(session.times * 1e6).astype(np.int64), se3_inverse(pose), vel,
                     np.zeros_like(vel), grav)


# ---- calibration files (projectaria MPS formats, read by SessionData::load, interfaces/ark/session_data/SessionData.cpp:81-273) -----
# Schemas restated from the readers the reference links (deps/projectaria_tools):
#   camera / IMU records   core/calibration/loader/SensorCalibrationJson.cpp:33-118 (parse), :185-244 (write)
#   SE3                    core/data_provider/json_io/JsonHelpers.h:77-96  {"Translation": [x, y, z], "UnitQuaternion": [w, [x, y, z]]}
#   factory_calibration.json   core/calibration/loader/DeviceCalibrationJson.cpp:70-153 (parse), :155-216 (write)
#   online_calibration.jsonl   core/mps/OnlineCalibrationsReader.cpp:27-83 (one JSON per line; :99-125 the pre-June-2023 variant whose
#                              CameraCalibrations / ImuCalibrations are Python-repr strings), :127-165
#   vrs_source_info.json       SessionData.cpp:84-103 {"camera_ids": [serial, ...], "imu_ids": [label, ...]}
#   reset_events.json          SessionData.cpp:337-358 {"reset_events": [{"tracking_timestamp_us": int}, ...]}
# Poses are rows (qx, qy, qz, qw, tx, ty, tz), the library's SE3 layout; IMU calibrations are the 32 doubles of a VAR_IMUCALIB
# variable (gyroScale3, accelScale3, accelBias3, gyroBias3, gyroNonorth9 and accelNonorth9 column-major, dtRefAccel, dtRefGyro:
# lib/motion/imu_types/ImuMeasurementModelParameters.h:28-60); camera intrinsics are the 15 Fisheye624 parameters.
import json

VRS_SOURCE_INFO_FILE = "vrs_source_info.json"
ONLINE_CALIBRATION_FILE = "online_calibration.jsonl"
FACTORY_CALIBRATION_FILE = "factory_calibration.json"
RESET_EVENTS_FILE = "reset_events.json"
FISHEYE624 = "FisheyeRadTanThinPrism"
# hard-coded Aria accelerometer noise (SessionData.cpp:211-224); any other label keeps the default noise model
IMU_ACCEL_SAMPLE_VARIANCE = {"imu-left": 7.7951241e-3, "imu-right": 6.6297049e-3}


def se3_to_json(T):
    T = np.asarray(T, np.float64)
    return {"Translation": [float(x) for x in T[4:7]], "UnitQuaternion": [float(T[3]), [float(x) for x in T[0:3]]]}


def se3_from_json(j):
    q = j["UnitQuaternion"]
    if len(q) != 2:
        raise ValueError(f"Expects UnitQuaternion to have two components, actual size: {len(q)}")
    return np.array([q[1][0], q[1][1], q[1][2], q[0]] + list(j["Translation"]), np.float64)


def se3_mul(A, B):
    """A * B for single poses (qx, qy, qz, qw, t)"""
    ax, ay, az, aw = A[:4]; bx, by, bz, bw = B[:4]
    q = np.array([aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx,
                  aw * bz + ax * by - ay * bx + az * bw, aw * bw - ax * bx - ay * by - az * bz])
    x, y, z, w = A[:4]
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    return np.concatenate([q, R @ B[4:7] + A[4:7]])


def se3_inv(T):
    return se3_inverse(np.asarray(T, np.float64)[None])[0]


def camera_calibration_to_json(cam):
    """cam: dict(label, serial, params[15], T_device_camera[7], width, height, max_solid_angle, valid_radius=None,
    time_offset=0.0) -> the record cameraCalibrationToJson writes (SensorCalibrationJson.cpp:185-209)"""
    j = {"Calibrated": True, "Label": cam["label"], "SerialNumber": cam["serial"], "T_Device_Camera": se3_to_json(cam["T_device_camera"]),
         "Projection": {"Description": "fx, cx, cy, k_0, k_1, k_2, k_3, k_4, k_5, p_0, p_1, s_0, s_1, s_2, s_3", "Name": FISHEYE624,
                        "Params": [float(x) for x in cam["params"]]}}
    if cam.get("time_offset", 0.0) != 0.0:
        j["TimeOffsetSec_Device_Camera"] = float(cam["time_offset"])
    j["ConfigData"] = {"ImageWidth": int(cam["width"]), "ImageHeight": int(cam["height"]), "MaxSolidAngle": float(cam["max_solid_angle"])}
    if cam.get("valid_radius") is not None:
        j["ConfigData"]["ValidRadius"] = float(cam["valid_radius"])
    return j


def camera_calibration_from_json(j):
    """parseCameraCalibrationFromJson (SensorCalibrationJson.cpp:33-88).  Only the Fisheye624 model is supported (the
    one ViBa optimises, CameraModelParam.h); a record without ConfigData needs the per-device sensor tables of
    projectaria's CameraConfigBuilder, which are not restated: ValueError."""
    if j["Projection"]["Name"] != FISHEYE624:
        raise ValueError(f"camera {j.get('Label')}: projection model {j['Projection']['Name']} is not supported")
    if "ConfigData" not in j:
        raise ValueError(f"No config found for camera {j['Label']} (a record without ConfigData)")
    c = j["ConfigData"]
    params = np.array(j["Projection"]["Params"], np.float64)
    if params.size != 15:
        raise ValueError(f"camera {j['Label']}: {params.size} projection parameters, Fisheye624 has 15")
    return {"label": j["Label"], "serial": j["SerialNumber"], "params": params, "T_device_camera": se3_from_json(j["T_Device_Camera"]),
            "width": int(c["ImageWidth"]), "height": int(c["ImageHeight"]), "max_solid_angle": float(c["MaxSolidAngle"]),
            "valid_radius": float(c["ValidRadius"]) if "ValidRadius" in c else None,
            "time_offset": float(j.get("TimeOffsetSec_Device_Camera", 0.0)), "readout_time": None}


def rescale_camera(cam, new_size, scale):
    """CameraCalibration::rescale (core/calibration/CameraCalibration.cpp:187-205) with Fisheye624::scaleParams
    (FisheyeRadTanThinPrism.h:243-256): f *= s, c = s (c + 0.5) - 0.5, valid radius *= s"""
    out = dict(cam)
    p = np.array(cam["params"], np.float64)
    p[0] *= scale
    p[1] = scale * (p[1] + 0.5) - 0.5
    p[2] = scale * (p[2] + 0.5) - 0.5
    out["params"] = p
    if cam.get("valid_radius") is not None:
        out["valid_radius"] = cam["valid_radius"] * scale
    out["width"], out["height"] = int(new_size[0]), int(new_size[1])
    return out


def imu_scale_matrices(calib32):
    """(gyro, accel) rectification matrices of a 32-double IMU calibration: diag(scale) * nonorth
    (ImuMeasurementModelParameters::get{Gyro,Accel}ScaleMat, :123-131)"""
    c = np.asarray(calib32, np.float64)
    gN, aN = c[12:21].reshape(3, 3).T, c[21:30].reshape(3, 3).T      # stored column-major
    return np.diag(c[0:3]) @ gN, np.diag(c[3:6]) @ aN


def imu_calibration_to_json(label, calib32, T_device_imu):
    """imuCalibrationToJson (SensorCalibrationJson.cpp:211-236) of toProjectAriaCalibration (ImuCalibConversion.cpp:25-38)"""
    c = np.asarray(calib32, np.float64)
    g, a = imu_scale_matrices(c)
    return {"Label": label, "SerialNumber": "", "Calibrated": True, "T_Device_Imu": se3_to_json(T_device_imu),
            "Accelerometer": {"Model": {"Name": "UpperTriagonalLinear", "RectificationMatrix": [[float(x) for x in r] for r in a]},
                              "Bias": {"Offset": [float(x) for x in c[6:9]], "Name": "Constant"}, "TimeOffsetSec_Device_Accel": float(c[30])},
            "Gyroscope": {"Bias": {"Name": "Constant", "Offset": [float(x) for x in c[9:12]]},
                          "Model": {"Name": "Linear", "RectificationMatrix": [[float(x) for x in r] for r in g]},
                          "TimeOffsetSec_Device_Gyro": float(c[31])}}


def imu_calibration_from_json(j):
    """parseImuCalibrationFromJson (SensorCalibrationJson.cpp:98-118) + fromProjectAriaCalibration
    (ImuCalibConversion.cpp:13-23) + setScaleMatrices (ImuMeasurementModelParameters.h:103-120: scale = row norms,
    nonorthogonality = normalised rows, the accelerometer's must be upper triangular): (label, calib32, T_device_imu)"""
    a = np.array(j["Accelerometer"]["Model"]["RectificationMatrix"], np.float64).reshape(3, 3)
    g = np.array(j["Gyroscope"]["Model"]["RectificationMatrix"], np.float64).reshape(3, 3)
    gs, as_ = np.linalg.norm(g, axis=1), np.linalg.norm(a, axis=1)
    gN, aN = g / gs[:, None], a / as_[:, None]
    if abs(aN[1, 0]) > 1e-14 or abs(aN[2, 0]) > 1e-14 or abs(aN[2, 1]) > 1e-14:
        raise ValueError("ImuMeasurementModelParameters::setScaleMatrices: accel should be upper triangular")
    c = np.concatenate([gs, as_, np.array(j["Accelerometer"]["Bias"]["Offset"], np.float64), np.array(j["Gyroscope"]["Bias"]["Offset"], np.float64),
                        gN.T.ravel(), aN.T.ravel(),
                        [float(j["Accelerometer"]["TimeOffsetSec_Device_Accel"]), float(j["Gyroscope"]["TimeOffsetSec_Device_Gyro"])]])
    return j["Label"], c, se3_from_json(j["T_Device_Imu"])


def write_factory_calibration(path, cameras, imus, origin_label, device_class="Aria", build_version=""):
    """deviceCalibrationToJson (DeviceCalibrationJson.cpp:155-216): cameras = camera dicts, imus = [(label, calib32,
    T_device_imu)]; the records are written sorted by label (std::map order)"""
    j = {"AlgorithmName": "RealCalib", "AlgorithmVersion": "1.0 Beta", "CalibrationSource": "Unknown", "Serial": "",
         "FileFormat": {"Tags": [], "Timestamp": "", "UnixTime": 0, "Version": 3},
         "DeviceClassInfo": {"BuildVersion": build_version, "DeviceClass": device_class},
         "OriginSpecification": {"Type": "Custom", "ChildLabel": origin_label}}
    if cameras:
        j["CameraCalibrations"] = [camera_calibration_to_json(c) for c in sorted(cameras, key=lambda c: c["label"])]
    if imus:
        j["ImuCalibrations"] = [imu_calibration_to_json(*i) for i in sorted(imus, key=lambda i: i[0])]
    with open(path, "w") as f:
        json.dump(j, f, indent=2)


def read_factory_calibration(path):
    """deviceCalibrationFromJson (DeviceCalibrationJson.cpp:70-153), cameras and IMUs only (magnetometer / barometer /
    microphone records are not used by ViBa): dict(cameras {label: cam}, imus {label: (calib32, T_device_imu)},
    origin_label, device_class, build_version)"""
    if not os.path.exists(path):
        raise FileNotFoundError(f"File not found: {path}")
    with open(path) as f:
        j = json.load(f)
    cams = {c["Label"]: camera_calibration_from_json(c) for c in j.get("CameraCalibrations", [])}
    imus = {}
    for i in j.get("ImuCalibrations", []):
        label, c, T = imu_calibration_from_json(i)
        imus[label] = (c, T)
    info = j.get("DeviceClassInfo", {})
    return {"cameras": dict(sorted(cams.items())), "imus": dict(sorted(imus.items())), "origin_label": j["OriginSpecification"]["ChildLabel"],
            "device_class": info.get("DeviceClass", "NotValid"), "build_version": info.get("BuildVersion", "")}


def write_online_calibration(path, records):
    """online_calibration.jsonl in the format written since June 2023: records = [dict(tracking_timestamp_us,
    utc_timestamp_ns, cameras [cam dicts, `readout_time` optional], imus [(label, calib32, T_device_imu)])]"""
    with open(path, "w") as f:
        for r in records:
            j = {"tracking_timestamp_us": int(r["tracking_timestamp_us"]), "utc_timestamp_ns": int(r["utc_timestamp_ns"]),
                 "CameraCalibrations": [camera_calibration_to_json(c) for c in r["cameras"]],
                 "ImuCalibrations": [imu_calibration_to_json(*i) for i in r["imus"]],
                 "ImageSizes": [[int(c["width"]), int(c["height"])] for c in r["cameras"]],
                 "ReadoutTimesSec": [[k, float(c["readout_time"])] for k, c in enumerate(r["cameras"]) if c.get("readout_time") is not None]}
            f.write(json.dumps(j) + "\n")


def _fix_json_string(text):   # fixJsonString (OnlineCalibrationsReader.cpp:93-99)
    return text.replace("True", "true").replace("False", "false").replace("'", '"')


def read_online_calibration(path):
    """readOnlineCalibration (OnlineCalibrationsReader.cpp:127-165): one record per line, both format generations; image
    size and readout time override the parsed record when the line carries them (:46-77).  A missing file gives [] as
    in the reference (SessionData::load then throws "Unable to load online calib!")."""
    if not os.path.exists(path):
        return []
    out = []
    with open(path) as f:
        for line in f:
            if not line.strip():
                continue
            doc = json.loads(line)
            in_str = isinstance(doc["tracking_timestamp_us"], str)
            cams_j = json.loads(_fix_json_string(doc["CameraCalibrations"])) if in_str else doc["CameraCalibrations"]
            imus_j = json.loads(_fix_json_string(doc["ImuCalibrations"])) if in_str else doc["ImuCalibrations"]
            cams = []
            for k, cj in enumerate(cams_j):
                cam = camera_calibration_from_json(cj)
                if not in_str and "ImageSizes" in doc:
                    cam["width"], cam["height"] = int(doc["ImageSizes"][k][0]), int(doc["ImageSizes"][k][1])
                    for idx, t in doc.get("ReadoutTimesSec", []):
                        if int(idx) == k:
                            cam["readout_time"] = float(t)
                            break
                cams.append(cam)
            out.append({"tracking_timestamp_us": int(doc["tracking_timestamp_us"]), "utc_timestamp_ns": int(doc["utc_timestamp_ns"]),
                        "cameras": cams, "imus": [imu_calibration_from_json(i) for i in imus_j]})
    return out


def write_vrs_source_info(path, camera_serials, imu_labels):
    with open(path, "w") as f:
        json.dump({"camera_ids": list(camera_serials), "imu_ids": list(imu_labels)}, f)


def read_vrs_source_info(path):
    if not os.path.exists(path):
        raise FileNotFoundError(f"File not found: {path}")
    with open(path) as f:
        j = json.load(f)
    return list(j["camera_ids"]), list(j["imu_ids"])


def write_reset_events(path, timestamps_us):
    with open(path, "w") as f:
        json.dump({"reset_events": [{"tracking_timestamp_us": int(t)} for t in timestamps_us]}, f)


def read_reset_events(path):
    """SessionData.cpp:337-358: optional file; absent -> no reset events"""
    if not os.path.exists(path):
        return []
    with open(path) as f:
        j = json.load(f)
    if not isinstance(j.get("reset_events"), list):
        raise ValueError("reset_events.json: no reset_events array")
    out = []
    for e in j["reset_events"]:
        if not isinstance(e, dict) or not isinstance(e.get("tracking_timestamp_us"), int):
            raise ValueError("reset_events.json: every event needs an integer tracking_timestamp_us")
        out.append(e["tracking_timestamp_us"])
    return out


def load_session_calibration(directory):
    """The calibration half of SessionData::load (interfaces/ark/session_data/SessionData.cpp:81-273): SLAM source info,
    factory calibration (cameras adapted to the online resolution, valid radius / time offset / readout time taken from the
    first online record, :159-186; extrinsics re-expressed against the body IMU = SLAM IMU n.0, :141-149,193-194,207-208;
    hard-coded accelerometer noise, :211-224) and every online calibration record (:236-273).  Same failures as the
    reference, as RuntimeError / FileNotFoundError with its texts."""
    cam_serials, imu_labels = read_vrs_source_info(os.path.join(directory, VRS_SOURCE_INFO_FILE))
    online = read_online_calibration(os.path.join(directory, ONLINE_CALIBRATION_FILE))
    if not online:
        raise RuntimeError("Unable to load online calib!")
    online_cam_index = {c["label"]: i for i, c in enumerate(online[0]["cameras"])}
    factory = read_factory_calibration(os.path.join(directory, FACTORY_CALIBRATION_FILE))
    sensors = {**{k: v["T_device_camera"] for k, v in factory["cameras"].items()}, **{k: v[1] for k, v in factory["imus"].items()}}
    if not imu_labels or imu_labels[0] not in sensors:
        raise RuntimeError(f"Slam's IMU n.0 = {imu_labels[0] if imu_labels else '<none>'} not present in factory calibration")
    T_device_bodyImu = sensors[imu_labels[0]]
    T_bodyImu_device = se3_inv(T_device_bodyImu)
    fac = {"imu_labels": list(factory["imus"]), "camera_labels": list(factory["cameras"]), "camera_serials": [], "cameras": [],
           "T_cam_bodyImu": [], "imu_calib": [], "T_imu_bodyImu": [], "imu_accel_sample_variance": []}
    for label in fac["camera_labels"]:
        cam = factory["cameras"][label]
        if label in online_cam_index:
            oc = online[0]["cameras"][online_cam_index[label]]
            adapted = cam
            if (oc["width"], oc["height"]) != (cam["width"], cam["height"]):
                adapted = rescale_camera(cam, (oc["width"], oc["height"]), float(oc["width"]) / cam["width"])
            cam = dict(adapted, valid_radius=oc["valid_radius"], time_offset=oc["time_offset"], readout_time=oc["readout_time"])
        fac["camera_serials"].append(cam["serial"])
        fac["cameras"].append(cam)
        fac["T_cam_bodyImu"].append(se3_mul(se3_inv(cam["T_device_camera"]), T_device_bodyImu))
    for label in fac["imu_labels"]:
        c, T_device_imu = factory["imus"][label]
        fac["imu_calib"].append(c)
        fac["T_imu_bodyImu"].append(se3_mul(se3_inv(T_device_imu), T_device_bodyImu))
        fac["imu_accel_sample_variance"].append(IMU_ACCEL_SAMPLE_VARIANCE.get(label))
    out_online = {"calibs": []}
    for rec in online:
        serials = [c["serial"] for c in rec["cameras"]]
        cam_labels = [c["label"] for c in rec["cameras"]]
        i_labels = [i[0] for i in rec["imus"]]
        out_online["calibs"].append({
            "timestamp_us": rec["tracking_timestamp_us"], "cameras": rec["cameras"],
            "T_cam_bodyImu": [se3_inv(se3_mul(T_bodyImu_device, c["T_device_camera"])) for c in rec["cameras"]],
            "imu_calib": [i[1] for i in rec["imus"]],
            "T_imu_bodyImu": [se3_inv(se3_mul(T_bodyImu_device, i[2])) for i in rec["imus"]]})
        if len(out_online["calibs"]) == 1:
            out_online.update(camera_serials=serials, camera_labels=cam_labels, imu_labels=i_labels)
        elif serials != out_online["camera_serials"] or cam_labels != out_online["camera_labels"] or i_labels != out_online["imu_labels"]:
            raise RuntimeError("mismatch in labels/serials")
    return {"slam_info": {"camera_serials": cam_serials, "imu_labels": imu_labels, "T_bodyImu_device": T_bodyImu_device},
            "factory": fac, "online": out_online,
            "reset_timestamps_us": read_reset_events(os.path.join(directory, RESET_EVENTS_FILE))}


SLAM_IMAGE_SIZE, RGB_IMAGE_SIZE = (640, 480), (1408, 1408)
CAMERA_LABELS = ["camera-slam-left", "camera-slam-right", "camera-rgb"]
IMU_LABELS = ["imu-left", "imu-right"]


def session_calibration(session, var_data=None):
    """(factory cameras, factory imus, online records) of a synthetic session in the file-level representation above.  The
    device frame is the body IMU's (T_Device_Imu of IMU 0 = identity); the variables hold T_Cam_BodyImu / T_Imu_BodyImu, the
    files T_Device_Camera / T_Device_Imu = their inverses; one online record per calibration window, stamped with the
    window's first rig; the factory file holds window 0."""
    g = session.graph
    data = g.var_data if var_data is None else var_data
    val = lambda v, n: np.array(data[g.var_off[v]:g.var_off[v] + n], np.float64)
    extr, intr = np.asarray(session.ids["extr"]), np.asarray(session.ids["intr"])
    n_win = extr.shape[1]
    ext = getattr(session, "ext", {}) or {}
    ident = np.array([0, 0, 0, 1, 0, 0, 0], np.float64)
    first_rig = [int(np.flatnonzero(session.win == w)[0]) if np.any(session.win == w) else 0 for w in range(n_win)]

    def cameras(w):
        out = []
        for c in range(extr.shape[0]):
            out.append({"label": CAMERA_LABELS[c], "serial": f"synthetic-slam-{c}", "params": val(intr[c][w], 15),
                        "T_device_camera": se3_inv(val(extr[c][w], 7)), "width": SLAM_IMAGE_SIZE[0], "height": SLAM_IMAGE_SIZE[1],
                        "max_solid_angle": 1.7, "valid_radius": None, "time_offset": 0.0, "readout_time": None})
        if "v_extr_rgb" in ext:
            cam = val(ext["v_intr_rgb"][w], 17)
            out.append({"label": CAMERA_LABELS[2], "serial": "synthetic-rgb", "params": cam[:15], "T_device_camera": se3_inv(val(ext["v_extr_rgb"][w], 7)),
                        "width": RGB_IMAGE_SIZE[0], "height": RGB_IMAGE_SIZE[1], "max_solid_angle": 1.1, "valid_radius": None,
                        "time_offset": float(cam[16]), "readout_time": float(cam[15])})
        return out

    def imus(w):
        out = [(IMU_LABELS[0], val(session.ids["imu"][w], 32), ident)]
        if "v_imu1" in ext:
            out.append((IMU_LABELS[1], val(ext["v_imu1"][w], 32), se3_inv(val(ext["v_imu_extr"][w], 7))))
        return out

    t_us = np.round(session.times * 1e6).astype(np.int64)
    online = [{"tracking_timestamp_us": int(t_us[first_rig[w]]), "utc_timestamp_ns": int(t_us[first_rig[w]]) * 1000,
               "cameras": cameras(w), "imus": imus(w)} for w in range(n_win)]
    return cameras(0), imus(0), online


def write_session_calibration(directory, session, var_data=None, imu_ids=None):
    """vrs_source_info.json, factory_calibration.json and online_calibration.jsonl of a synthetic session; imu_ids = the SLAM
    IMU labels of vrs_source_info.json (default: every IMU of the session; SessionData::load reads one sample file each)"""
    os.makedirs(directory, exist_ok=True)
    cams, imus, online = session_calibration(session, var_data)
    write_vrs_source_info(os.path.join(directory, VRS_SOURCE_INFO_FILE), [c["serial"] for c in cams[:2]],
                          [i[0] for i in imus] if imu_ids is None else list(imu_ids))
    write_factory_calibration(os.path.join(directory, FACTORY_CALIBRATION_FILE), cams, imus, IMU_LABELS[0])
    write_online_calibration(os.path.join(directory, ONLINE_CALIBRATION_FILE), online)
