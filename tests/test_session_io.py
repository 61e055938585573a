"""Session files in the reference's formats: the column lists are the reference's
(PointObservationFormat.h:13-23, ImuDataFormat.h:14-23, SaveDeviceTrajectory.cpp:16-37), values survive a write / read
round trip at the precision the reference stores them in (float32 observations and IMU samples), and the observation
table reproduces the visual stores of the session it came from."""
import numpy as np
from visual_inertial_bundle_adjustment_b200 import synth, session_io as sio, graph as G


def test_round_trip(tmp_path):
    s = synth.make_session(**synth.CONFIGS["tiny_rs"])
    sio.write_session(tmp_path, s)
    with open(tmp_path / sio.OBSERVATIONS_FILE) as f:
        assert f.readline().strip().split(",") == sio.OBSERVATION_COLUMNS and len(sio.OBSERVATION_COLUMNS) == 9
    with open(tmp_path / sio.IMU_FILE_PATTERN.format(0)) as f:
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
    t, g, a = sio.read_imu(tmp_path / sio.IMU_FILE_PATTERN.format(0))
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
