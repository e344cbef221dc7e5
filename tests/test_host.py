"""The C++ host layer on the CPU: config reader, ray generator, mesh readers, PointCloud2 header, trajectory.
Mirrors SRC/test/LidarDevice_test.cpp and the load checks of SRC/test/EmbreeTracer_test.cpp:86-91."""
import ctypes
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, xt32_sensor_args
from lidarshooter_b200 import hostapi, synth
from oracle import lso
from test_abi import declared_symbols

CFG0 = os.path.join(GOLDEN, "xt32_lidar_0000.json")
CFG1 = os.path.join(GOLDEN, "xt32_lidar_0001.json")


def test_host_library_exports_every_declared_symbol():
    names = [n for n in declared_symbols("lsb200_host.h") if n.startswith("lsbh_")]
    assert sorted(hostapi.SYMBOLS) == names
    L = ctypes.CDLL(hostapi.LIB_PATH)
    for n in names:
        assert hasattr(L, n), n


def test_lidar_device_config():
    """SRC/test/LidarDevice_test.cpp:51-59"""
    d = hostapi.LidarDevice(CFG0)
    assert d.sensor_uid == "lidar_0000"
    state, _, vi, hi = d.next_ray()
    d.reset()
    assert d.total_rays == 150 * 32 and d.total_channels == 32 and d.scan_ray_count == 150
    v, b, e, step = d.channels()
    assert np.array_equal(v, np.arange(15, -17, -1, dtype=np.float32)) and b == 0 and e == 360
    assert step == lso.azimuth_step(0.0, 360.0, 150)
    d2 = hostapi.LidarDevice(CFG0, "other_uid")
    assert d2.sensor_uid == "other_uid"


def test_lidar_device_limits_from_config():
    """detection_radius / horizontal_angle_range (config/hesai-pandar-XT-32-lidar_0000.json:14-18)"""
    dev = hostapi.LidarDevice(os.path.join(GOLDEN, "xt32_lidar_0000.json"))
    assert dev.limits() == (1000.0, -180.0, 180.0)
    dev.close()


def test_lidar_device_init_message():
    """SRC/test/LidarDevice_test.cpp:61-76"""
    d = hostapi.LidarDevice(CFG0)
    h, frame_id, fields = d.init_message(100)
    assert (h.n_fields, h.height, h.width, h.point_step, h.row_step) == (5, 1, 0, 32, 0)
    assert h.is_bigendian == 0 and h.is_dense == 1 and h.seq == 100
    assert frame_id == "PandarXT-32"
    assert fields == [("x", 0, 7, 1), ("y", 4, 7, 1), ("z", 8, 7, 1), ("intensity", 16, 7, 1), ("ring", 20, 4, 1)]


@pytest.mark.parametrize("uid", ["0000", "0001"])
def test_sensor_transform_and_rays_match_oracle(uid):
    d = hostapi.LidarDevice(os.path.join(GOLDEN, f"xt32_lidar_{uid}.json"))
    s = lso.Sensor(*xt32_sensor_args(uid))
    q, t, R, Ri = d.transform()
    assert np.array_equal(q, s.q) and np.array_equal(t, s.T)
    assert np.array_equal(R, s.R) and np.array_equal(Ri, s.Rinv)
    dirs = d.all_rays()
    assert np.array_equal(dirs, lso.ray_dirs(s.tables))
    # the stateful generator walks the same order, horizontal fastest, and reports the wrap
    d.reset()
    for k in range(151):
        state, dd, vi, hi = d.next_ray()
        assert np.array_equal(dd, dirs[k]) and state == 0
    assert (vi, hi) == (1, 1)
    p = np.array([3.0, -2.0, 1.5], np.float32)
    ps = d.origin_to_sensor(p)
    assert np.array_equal(ps, lso.transform_vertices(p[None], lso.IDENTITY_AFFINE, s.Rinv, s.T)[0])
    assert np.allclose(d.origin_to_sensor(ps, inverse=True), p, atol=1e-5)


def test_wrap_state_of_ray_generator():
    d = hostapi.LidarDevice(synthetic=(np.array([1.0, -1.0], np.float32), 0.0, 90.0, 3, (1, 0, 0, 0), (0, 0, 0)))
    states = [d.next_ray()[0] for _ in range(6)]
    assert states == [0, 0, 0, 0, 0, 1]            # SRC/LidarDevice.cpp:824-845


def test_transform_file_precedence(tmp_path):
    """inline sensorConfig > sensorConfigFile (SRC/LidarDevice.cpp:504-531); a missing file raises"""
    from make_golden import load_json
    cfg = load_json(CFG0)
    inline = cfg["device"].pop("sensorConfig")
    tf = tmp_path / "transform.json"
    tf.write_text("// a comment, like the reference's files have\n" + json.dumps(inline))
    cfg["device"]["sensorConfigFile"] = str(tf)
    p = tmp_path / "dev.json"
    p.write_text(json.dumps(cfg))
    d = hostapi.LidarDevice(str(p))
    assert np.array_equal(d.transform()[0], hostapi.LidarDevice(CFG0).transform()[0])
    cfg["device"]["sensorConfigFile"] = str(tmp_path / "missing.json")
    p.write_text(json.dumps(cfg))
    with pytest.raises(RuntimeError, match="File not found"):
        hostapi.LidarDevice(str(p))
    with pytest.raises(RuntimeError):
        hostapi.LidarDevice(str(tmp_path / "nope.json"))


def test_mesh_readers(meshes):
    """ground.stl -> 98 vertices, 162 triangles (SRC/test/EmbreeTracer_test.cpp:86-91)"""
    m = hostapi.Mesh(os.path.join(GOLDEN, "ground.stl"))
    nv, nt, step = m.counts()
    assert (nv, nt, step) == (98, 162, 16)
    v, t = m.arrays()
    assert np.array_equal(v, meshes["ground_v"]) and np.array_equal(t, meshes["ground_t"])
    o = hostapi.Mesh(os.path.join(GOLDEN, "ben_head.obj"))
    nv, nt, _ = o.counts()
    assert nv > 0 and nt > 0
    ov, ot = o.arrays()
    assert ot.max() < nv
    with pytest.raises(RuntimeError):
        hostapi.Mesh(os.path.join(GOLDEN, "does_not_exist.stl"))


def test_trajectory_player():
    """config/trajectory.json: 10 s at v=(0,5,0), w=(0,0,1) then 5 s at v=(1,0,0), w=(0,0,-1) -> 150 ticks of 0.1 s"""
    poses = hostapi.trajectory_poses(os.path.join(GOLDEN, "trajectory.json"), 0.1)
    assert poses.shape == (150, 6)
    ref = synth.trajectory_poses([(10.0, (0, 5, 0), (0, 0, 1)), (5.0, (1, 0, 0), (0, 0, -1))], 0.1)
    ref = np.array([np.concatenate(p) for p in ref], np.float32)
    assert np.allclose(poses, ref, rtol=0, atol=2e-4)
    assert np.allclose(poses[99, 3:], [0, 0, 10.0], atol=1e-4) and np.allclose(poses[-1, 3:], [0, 0, 5.0], atol=1e-4)


def _ros1_pointcloud2(seq, sec, nsec, frame_id, height, width, fields, bigendian, point_step, row_step, data, dense):
    """sensor_msgs/PointCloud2 as roscpp serialises it (little-endian, length-prefixed strings and arrays)."""
    import struct
    out = struct.pack("<III", seq, sec, nsec) + struct.pack("<I", len(frame_id)) + frame_id.encode()
    out += struct.pack("<II", height, width) + struct.pack("<I", len(fields))
    for name, offset, datatype, count in fields:
        out += struct.pack("<I", len(name)) + name.encode() + struct.pack("<IBI", offset, datatype, count)
    out += struct.pack("<BII", int(bigendian), point_step, row_step) + struct.pack("<I", len(data)) + bytes(data)
    return out + struct.pack("<B", int(dense))


XT32_FIELDS = [("x", 0, 7, 1), ("y", 4, 7, 1), ("z", 8, 7, 1), ("intensity", 16, 7, 1), ("ring", 20, 4, 1)]


def test_pointcloud2_wire_reader_round_trip():
    """host/CloudSerializer: parse + serialise of a hand-packed ROS-1 message is the identity; truncation is refused."""
    rng = np.random.default_rng(3)
    data = rng.integers(0, 256, 7 * 32, dtype=np.uint8)
    msg = _ros1_pointcloud2(41, 1700000000, 123456789, "PandarXT-32", 1, 7, XT32_FIELDS, False, 32, 0, data, True)
    again, h, stamp = hostapi.reserialize_cloud(np.frombuffer(msg, np.uint8))
    assert bytes(again) == msg
    assert (h.seq, h.height, h.width, h.point_step, h.row_step, h.n_fields, h.is_dense) == (41, 1, 7, 32, 0, 5, 1)
    assert abs(stamp - 1700000000.123456789) < 1e-6
    empty = _ros1_pointcloud2(0, 0, 0, "", 1, 0, [], False, 32, 0, b"", False)
    assert bytes(hostapi.reserialize_cloud(np.frombuffer(empty, np.uint8))[0]) == empty
    for cut in (3, 20, len(msg) - 1, len(msg) - 40):
        with pytest.raises(ValueError):
            hostapi.reserialize_cloud(np.frombuffer(msg[:cut], np.uint8))


def test_reachable_nodes_masks_records_no_reference_leads_to():
    """capi.reachable_nodes (tree comparisons of the GPU tests): rows reached from the root through child references >= 0
    (minus the root offset) are kept, every other row is zeroed; negative references are leaves."""
    import numpy as np
    from lidarshooter_b200 import capi
    root = 1024
    nodes = np.full((6, 16), 7.0, np.float32)
    child = nodes[:, 12:14].view(np.int32)
    child[0] = (root + 1, root + 4)          # root -> nodes 1 and 4
    child[1] = (~3, ~(5 | (1 << 28)))        # two leaf references
    child[4] = (root + 5, ~9)
    child[5] = (~1, ~2)
    child[2] = (root + 3, root + 3)          # rows 2 and 3: garbage nobody references
    child[3] = (~0, ~0)
    out = capi.reachable_nodes(nodes, root)
    assert np.array_equal(out[[0, 1, 4, 5]].view(np.uint8), nodes[[0, 1, 4, 5]].view(np.uint8))      # bytes: leaf references are NaN patterns
    assert not out[[2, 3]].any()
    assert nodes[2].any()                    # the input is not modified
