"""Session files in the reference's formats: the column lists are the reference's
(PointObservationFormat.h:13-23, ImuDataFormat.h:14-23, SaveDeviceTrajectory.cpp:16-37), values survive a write / read
round trip at the precision the reference stores them in (float32 observations and IMU samples), and the observation
table reproduces the visual stores of the session it came from."""
import numpy as np
import pytest
from visual_inertial_bundle_adjustment_b200 import synth, session_io as sio, graph as G


def test_round_trip(tmp_path):
    s = synth.make_session(**synth.CONFIGS["tiny_rs"])
    sio.write_session(tmp_path, s)
    with open(tmp_path / sio.OBSERVATIONS_FILE) as f:
        assert f.readline().strip().split(",") == sio.OBSERVATION_COLUMNS and len(sio.OBSERVATION_COLUMNS) == 9
    with open(tmp_path / sio.IMU_FILE_PATTERN.format(sio.IMU_LABELS[0])) as f:
        assert f.readline().strip().split(",") == sio.IMU_COLUMNS and len(sio.IMU_COLUMNS) == 8
    with open(tmp_path / sio.TRAJECTORY_FILE) as f:
        assert f.readline().strip().split(",") == sio.TRAJECTORY_COLUMNS and len(sio.TRAJECTORY_COLUMNS) == 20
    pid, t_ns, cam, uv, sh = sio.read_observations(tmp_path / sio.OBSERVATIONS_FILE)
    p0, t0, c0, uv0, sh0 = sio.session_observations(s)
    n_vis = sum(st.vars.shape[0] for st in s.graph.stores if st.type in (G.F_VISUAL, G.F_VISUAL_RS))
    assert pid.size == n_vis and np.array_equal(pid, p0) and np.array_equal(t_ns, t0) and np.array_equal(cam, c0)
    assert set(np.unique(cam)) == {0, 1, 2}
    assert np.array_equal(uv, uv0.astype(np.float32)) and np.array_equal(sh, sh0.astype(np.float32))
    # every capture timestamp is a rig timestamp (Matcher.cpp:33-60 joins on them)
    assert set(t_ns.tolist()) <= set(np.round(s.times * 1e9).astype(np.int64).tolist())
    t, g, a = sio.read_imu(tmp_path / sio.IMU_FILE_PATTERN.format(sio.IMU_LABELS[0]))
    tr, gr, ar = synth._imu0_raw(s)
    assert np.array_equal(t, tr) and np.array_equal(g, gr) and np.array_equal(a, ar) and np.all(np.diff(t) > 0)
    tj = sio.read_trajectory(tmp_path / sio.TRAJECTORY_FILE)
    assert np.array_equal(tj["t_us"], np.round(s.times * 1e6).astype(np.int64))
    gr_ = s.graph
    pose = np.stack([gr_.var_data[gr_.var_off[v]:gr_.var_off[v] + 7] for v in s.ids["pose"]])
    back = sio.se3_inverse(tj["T_world_device"])
    assert np.abs(back - pose).max() <= 1e-12 and np.abs(np.linalg.norm(tj["T_world_device"][:, :4], axis=1) - 1).max() <= 1e-12
    assert np.allclose(tj["gravity_world"], [0, 0, -9.81])


def test_reader_rejects_other_columns(tmp_path):
    p = tmp_path / "x.csv"
    p.write_text("a,b,c\n1,2,3\n")
    for fn in (sio.read_observations, sio.read_imu, sio.read_trajectory):
        try:
            fn(p)
            assert False
        except ValueError as ex:
            assert "unexpected columns" in str(ex)


# ---- calibration files (projectaria MPS JSON formats + the calibration half of SessionData::load) -------------------------
def _var(g, v, n):
    return g.var_data[g.var_off[v]:g.var_off[v] + n]


def test_calibration_files_round_trip(tmp_path):
    """a synthetic session's calibration written in the files SessionData::load reads (SessionData.cpp:29-31) comes back, through
    the reference's own conversions (device-frame extrinsics -> T_Cam_BodyImu / T_Imu_BodyImu, rectification matrices ->
    scale + nonorthogonality), as the values of the session's variables"""
    import json
    s = synth.make_session(**synth.CONFIGS["tiny_rs"])
    g = s.graph
    sio.write_session(tmp_path, s)
    for f in (sio.VRS_SOURCE_INFO_FILE, sio.FACTORY_CALIBRATION_FILE, sio.ONLINE_CALIBRATION_FILE, sio.OBSERVATIONS_FILE, sio.TRAJECTORY_FILE,
              sio.IMU_FILE_PATTERN.format("imu-left")):
        assert (tmp_path / f).exists(), f
    c = sio.load_session_calibration(tmp_path)
    assert c["slam_info"]["imu_labels"] == ["imu-left"] and len(c["slam_info"]["camera_serials"]) == 2
    assert np.array_equal(c["slam_info"]["T_bodyImu_device"][:4] != 0, [False, False, False, True])
    n_win = s.n_win
    assert len(c["online"]["calibs"]) == n_win
    assert c["online"]["camera_labels"] == sio.CAMERA_LABELS and c["online"]["imu_labels"] == sio.IMU_LABELS
    extr, intr = np.asarray(s.ids["extr"]), np.asarray(s.ids["intr"])
    for w, rec in enumerate(c["online"]["calibs"]):
        assert rec["timestamp_us"] == int(np.round(s.times[np.flatnonzero(s.win == w)[0]] * 1e6))
        for cam in range(2):
            assert np.abs(rec["T_cam_bodyImu"][cam] - _var(g, extr[cam][w], 7)).max() <= 1e-15
            assert np.array_equal(rec["cameras"][cam]["params"], _var(g, intr[cam][w], 15))
            assert (rec["cameras"][cam]["width"], rec["cameras"][cam]["height"]) == sio.SLAM_IMAGE_SIZE
            assert rec["cameras"][cam]["readout_time"] is None and rec["cameras"][cam]["time_offset"] == 0.0
        rgb = _var(g, s.ext["v_intr_rgb"][w], 17)
        assert np.array_equal(rec["cameras"][2]["params"], rgb[:15])
        assert rec["cameras"][2]["readout_time"] == rgb[15] and rec["cameras"][2]["time_offset"] == rgb[16]
        assert np.abs(rec["T_cam_bodyImu"][2] - _var(g, s.ext["v_extr_rgb"][w], 7)).max() <= 1e-15
        assert np.abs(rec["imu_calib"][0] - _var(g, s.ids["imu"][w], 32)).max() <= 1e-15
        assert np.abs(rec["imu_calib"][1] - _var(g, s.ext["v_imu1"][w], 32)).max() <= 1e-15
        assert np.abs(rec["T_imu_bodyImu"][0] - [0, 0, 0, 1, 0, 0, 0]).max() == 0
        assert np.abs(rec["T_imu_bodyImu"][1] - _var(g, s.ext["v_imu_extr"][w], 7)).max() <= 1e-15
    # factory: records in label order (std::map), window 0, the hard-coded Aria accelerometer noise (SessionData.cpp:211-224)
    fac = c["factory"]
    assert fac["camera_labels"] == sorted(sio.CAMERA_LABELS) and fac["imu_labels"] == sorted(sio.IMU_LABELS)
    assert fac["imu_accel_sample_variance"] == [7.7951241e-3, 6.6297049e-3]
    k = fac["camera_labels"].index("camera-slam-right")
    assert np.abs(fac["T_cam_bodyImu"][k] - _var(g, extr[1][0], 7)).max() <= 1e-15
    # the factory RGB record takes radius / time offset / readout time from the first online record (:175-186)
    k = fac["camera_labels"].index("camera-rgb")
    assert fac["cameras"][k]["readout_time"] == c["online"]["calibs"][0]["cameras"][2]["readout_time"]
    # the files are what the reference's parsers expect: SE3 as {"Translation", "UnitQuaternion": [w, [x, y, z]]}, upper
    # triangular accelerometer matrix, one JSON document per line
    with open(tmp_path / sio.ONLINE_CALIBRATION_FILE) as f:
        lines = f.read().strip().split("\n")
    assert len(lines) == n_win
    doc = json.loads(lines[0])
    assert set(doc) >= {"tracking_timestamp_us", "utc_timestamp_ns", "CameraCalibrations", "ImuCalibrations", "ImageSizes", "ReadoutTimesSec"}
    q = doc["CameraCalibrations"][0]["T_Device_Camera"]["UnitQuaternion"]
    assert len(q) == 2 and len(q[1]) == 3 and doc["CameraCalibrations"][0]["Projection"]["Name"] == "FisheyeRadTanThinPrism"
    a = np.array(doc["ImuCalibrations"][0]["Accelerometer"]["Model"]["RectificationMatrix"])
    assert a[1, 0] == 0 and a[2, 0] == 0 and a[2, 1] == 0
    assert doc["ReadoutTimesSec"] == [[2, doc["ReadoutTimesSec"][0][1]]]


def test_factory_camera_is_adapted_to_the_online_resolution(tmp_path):
    """SessionData.cpp:163-186 + CameraCalibration::rescale + Fisheye624::scaleParams: f *= s, c = s (c + 0.5) - 0.5"""
    s = synth.make_session(**synth.CONFIGS["tiny"])
    cams, imus, online = sio.session_calibration(s)
    big = [dict(c, width=1280, height=960, valid_radius=700.0) for c in cams]
    for c in big:
        p = c["params"].copy(); p[0] *= 2; p[1] = 2 * (p[1] + 0.5) - 0.5; p[2] = 2 * (p[2] + 0.5) - 0.5
        c["params"] = p
    online[0]["cameras"][0]["valid_radius"] = 333.0
    sio.write_vrs_source_info(tmp_path / sio.VRS_SOURCE_INFO_FILE, [c["serial"] for c in cams], ["imu-left"])
    sio.write_factory_calibration(tmp_path / sio.FACTORY_CALIBRATION_FILE, big, imus, "imu-left")
    sio.write_online_calibration(tmp_path / sio.ONLINE_CALIBRATION_FILE, online)
    c = sio.load_session_calibration(tmp_path)
    k = c["factory"]["camera_labels"].index("camera-slam-left")
    got = c["factory"]["cameras"][k]
    assert (got["width"], got["height"]) == (640, 480)
    assert np.abs(got["params"] - cams[0]["params"]).max() <= 1e-12
    assert got["valid_radius"] == 333.0          # the online record's, not the rescaled factory one


def test_old_online_calibration_format_and_failures(tmp_path):
    import json
    s = synth.make_session(**synth.CONFIGS["tiny"])
    cams, imus, online = sio.session_calibration(s)
    # before June 2023 the calibration content was a Python-repr string (OnlineCalibrationsReader.cpp:86-125)
    new = tmp_path / "new.jsonl"
    sio.write_online_calibration(new, online[:1])
    doc = json.loads(new.read_text())
    old = {"tracking_timestamp_us": str(doc["tracking_timestamp_us"]), "utc_timestamp_ns": str(doc["utc_timestamp_ns"]),
           "CameraCalibrations": repr(doc["CameraCalibrations"]), "ImuCalibrations": repr(doc["ImuCalibrations"])}
    assert "True" in old["CameraCalibrations"] and "'" in old["CameraCalibrations"]
    (tmp_path / "old.jsonl").write_text(json.dumps(old) + "\n")
    a, b = sio.read_online_calibration(new)[0], sio.read_online_calibration(tmp_path / "old.jsonl")[0]
    assert a["tracking_timestamp_us"] == b["tracking_timestamp_us"]
    assert np.array_equal(a["cameras"][1]["params"], b["cameras"][1]["params"]) and np.array_equal(a["imus"][0][1], b["imus"][0][1])
    # failures of SessionData::load, with its texts
    d = tmp_path / "sess"; d.mkdir()
    with pytest.raises(FileNotFoundError, match="File not found"):
        sio.load_session_calibration(d)
    sio.write_vrs_source_info(d / sio.VRS_SOURCE_INFO_FILE, ["a", "b"], ["imu-other"])
    with pytest.raises(RuntimeError, match="Unable to load online calib"):
        sio.load_session_calibration(d)
    sio.write_online_calibration(d / sio.ONLINE_CALIBRATION_FILE, online)
    sio.write_factory_calibration(d / sio.FACTORY_CALIBRATION_FILE, cams, imus, "imu-left")
    with pytest.raises(RuntimeError, match="Slam's IMU n.0 = imu-other not present in factory calibration"):
        sio.load_session_calibration(d)
    sio.write_vrs_source_info(d / sio.VRS_SOURCE_INFO_FILE, ["a", "b"], ["imu-left"])
    bad = [dict(r) for r in online]
    bad[-1] = dict(bad[-1], cameras=[dict(c, serial="other") for c in bad[-1]["cameras"]])
    sio.write_online_calibration(d / sio.ONLINE_CALIBRATION_FILE, bad)
    with pytest.raises(RuntimeError, match="mismatch in labels/serials"):
        sio.load_session_calibration(d)
    # setScaleMatrices: the accelerometer matrix must be upper triangular (ImuMeasurementModelParameters.h:112-119)
    j = sio.imu_calibration_to_json(*imus[0])
    j["Accelerometer"]["Model"]["RectificationMatrix"][2][0] = 1e-3
    with pytest.raises(ValueError, match="accel should be upper triangular"):
        sio.imu_calibration_from_json(j)
    # reset events (SessionData.cpp:337-358): optional file
    assert sio.read_reset_events(d / sio.RESET_EVENTS_FILE) == []
    sio.write_reset_events(d / sio.RESET_EVENTS_FILE, [10, 2000000])
    sio.write_online_calibration(d / sio.ONLINE_CALIBRATION_FILE, online)
    assert sio.load_session_calibration(d)["reset_timestamps_us"] == [10, 2000000]
