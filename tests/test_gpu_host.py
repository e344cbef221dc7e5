"""lidarshooter::B200Tracer (the ITracer drop-in) on the GPU: the reference's own tracer tests, re-expressed.
SRC/test/EmbreeTracer_test.cpp:93-135 and SRC/test/OptixTracer_test.cpp:93-311."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, xt32_sensor_args
from lidarshooter_b200 import hostapi
from oracle import lso

pytestmark = pytest.mark.gpu
TRIANGLE, QUAD, GRID = 0, 1, 2
IDENT = lso.IDENTITY_AFFINE


@pytest.fixture()
def setup(meshes):
    dev = hostapi.LidarDevice(os.path.join(GOLDEN, "xt32_lidar_0000.json"))
    tracer = hostapi.B200Tracer(dev)
    ground = hostapi.Mesh(os.path.join(GOLDEN, "ground.stl"))
    ben = hostapi.Mesh(arrays=(meshes["ben_v"], meshes["ben_t"]))
    yield dev, tracer, ground, ben
    tracer.close()


def test_geometry_bookkeeping(setup):
    dev, tracer, ground, ben = setup
    assert tracer.getGeometryCount() == 0
    assert tracer.addGeometry("mesh", TRIANGLE, 98, 162) == 0
    assert tracer.getGeometryCount() == 1                              # EmbreeTracer_test.cpp:93-97
    assert tracer.addGeometry("grid", GRID, 4, 1) == 0                 # unsupported type: "returns false"
    assert tracer.getGeometryCount() == 1
    gid = tracer.addGeometry("second", TRIANGLE, 2823, 5489)
    assert gid == 1 and tracer.getGeometryId("second") == 1
    assert tracer.removeGeometry("second") == 1                        # id round trip, EmbreeTracer_test.cpp:99-114
    assert tracer.removeGeometry("second") == -1
    assert tracer.getGeometryId("second") == -1
    assert tracer.updateGeometry("nope", ground) == -108               # TraceException code 8
    assert tracer.updateGeometry("mesh", None) == -301                 # MeshNotSetException


def test_trace_scene_cloud(setup, meshes):
    """EmbreeTracer_test.cpp:122-135 (1668) and OptixTracer_test.cpp:122-311 (1781 -> 1668 -> 0)"""
    dev, tracer, ground, ben = setup
    tracer.setFlags(ordered=True)
    assert tracer.addGeometry("ground", TRIANGLE, 98, 162) == 0
    assert tracer.updateGeometry("ground", ground, IDENT) == 0
    assert tracer.commitScene() == 0
    assert tracer.traceScene(0) == 0
    h, pts = tracer.getTraceCloud()
    assert h.width == 1668 and pts.shape == (1668, 32) and h.height == 1 and h.point_step == 32 and h.seq == 0

    assert tracer.addGeometry("ben", TRIANGLE, 2823, 5489) == 1
    assert tracer.updateGeometry("ben", ben, translation=(0, 0, 0), rotation=(0, 0, 0)) == 0
    assert tracer.updateGeometry("ground", ground, IDENT) == 0
    assert tracer.commitScene() == 0
    assert tracer.traceScene(1) == 0
    h, pts = tracer.getTraceCloud()
    assert h.width == 1781 and h.seq == 1
    s = lso.Sensor(*xt32_sensor_args("0000"))
    ms = [(meshes["ground_v"], meshes["ground_t"], IDENT), (meshes["ben_v"], meshes["ben_t"], IDENT)]
    # default cloud bytes == the Embree path's: ring = 0 for every point (SRC/EmbreeTracer.cpp:333-334)
    ref0 = lso.scan(s, ms, ring_mode=1)
    assert np.array_equal(pts, ref0["points"])
    assert (pts.view(np.int32).reshape(-1, 8)[:, 5] == 0).all()
    # opt-in: ring = channel index, what the OptiX path writes (SRC/LidarDeviceKernels.cu:51)
    tracer.setFlags(ordered=True, ring_zero=False)
    assert tracer.traceScene(1) == 0
    h, pts = tracer.getTraceCloud()
    ref = lso.scan(s, ms)
    assert np.array_equal(pts, ref["points"])
    assert set(np.unique(pts.view(np.int32).reshape(-1, 8)[:, 5])) <= set(range(32)) and pts.view(np.int32).reshape(-1, 8)[:, 5].max() > 0
    # the spdlog logger given to create() (SRC/EmbreeTracer.hpp:51) received the library's lines
    lines = tracer.logLines()
    assert any(lv == 2 and "lidar_0000" in m and "4800 rays" in m for lv, m in lines), lines
    assert tracer.removeGeometry("no-such-mesh") == -1
    lines = tracer.logLines()
    assert lines[-1][0] == 4 and "no-such-mesh" in lines[-1][1]
    rc, t, tri = tracer.traceRaw()
    assert rc == 0 and np.array_equal(tri, ref["tri"]) and np.array_equal(t, ref["t"])

    # the XT-32 config's own limits (1000 m, +-180 deg) change nothing
    assert tracer.setSensorLimits(True) == 0 and tracer.traceScene(1) == 0 and tracer.getTraceCloud()[0].width == 1781
    assert tracer.setSensorLimits(False) == 0

    # ROS-1 wire bytes of the cloud: header of initMessage + the 32-byte records, row_step 0 like the
    # reference or width * point_step behind the flag
    from test_host import _ros1_pointcloud2, XT32_FIELDS
    for fix, row_step in ((False, 0), (True, 1781 * 32)):
        wire = tracer.serializeCloud(stamp=12.25, fix_row_step=fix)
        want = _ros1_pointcloud2(1, 12, 250000000, "PandarXT-32", 1, 1781, XT32_FIELDS, False, 32, row_step, pts.tobytes(), True)
        assert bytes(wire) == want
    path = os.path.join(os.environ.get("TMPDIR", "/tmp"), "lsb200_cloud_stream.bin")
    if os.path.exists(path):
        os.remove(path)
    assert tracer.appendCloudToFile(path) == 0 and tracer.appendCloudToFile(path, fix_row_step=True) == 0
    blob = open(path, "rb").read()
    n0 = int.from_bytes(blob[:4], "little")
    assert blob[4:4 + n0] == _ros1_pointcloud2(1, 12, 250000000, "PandarXT-32", 1, 1781, XT32_FIELDS, False, 32, 0, pts.tobytes(), True)
    assert int.from_bytes(blob[4 + n0:8 + n0], "little") == len(blob) - 8 - n0
    os.remove(path)

    # world-frame points = CloudTransformer::applyInverseTransform (SRC/CloudTransformer.cpp:202-265), on the device
    tracer.setFlags(ordered=True, ring_zero=False, world_points=True)
    assert tracer.traceScene(2) == 0
    _, wpts = tracer.getTraceCloud()
    sp = ref["points"].view(np.float32).reshape(-1, 8)[:, :3]
    expect = np.stack([dev.origin_to_sensor(p, inverse=True) for p in sp])
    assert np.array_equal(wpts.view(np.float32).reshape(-1, 8)[:, :3], expect)
    tracer.setFlags(ordered=True, ring_zero=False)

    assert tracer.removeGeometry("ben") == 1
    assert tracer.getGeometryCount() == 1
    assert tracer.traceScene(3) == 0
    assert tracer.getTraceCloud()[0].width == 1668
    assert tracer.removeGeometry("ground") == 0
    assert tracer.getGeometryCount() == 0
    assert tracer.commitScene() == -1
    assert tracer.traceScene(4) == -1
    h, pts = tracer.getTraceCloud()
    assert h.width == 0 and pts.shape[0] == 0 and h.point_step == 32 and h.seq == 4


def test_trajectory_frames_two_devices(meshes):
    """BASELINE config 2: lidar_0000 + lidar_0001, ben follows config/trajectory.json, ground static, refit path."""
    poses = hostapi.trajectory_poses(os.path.join(GOLDEN, "trajectory.json"), 0.1)
    ground = hostapi.Mesh(os.path.join(GOLDEN, "ground.stl"))
    ben = hostapi.Mesh(arrays=(meshes["ben_v"], meshes["ben_t"]))
    for uid in ("0000", "0001"):
        dev = hostapi.LidarDevice(os.path.join(GOLDEN, f"xt32_lidar_{uid}.json"))
        tracer = hostapi.B200Tracer(dev)
        tracer.setFlags(ordered=True)
        tracer.addGeometry("ground", TRIANGLE, 98, 162)
        tracer.addGeometry("ben", TRIANGLE, 2823, 5489)
        s = lso.Sensor(*xt32_sensor_args(uid))
        for k in range(0, 150, 7):
            lin, ang = poses[k, :3], poses[k, 3:]
            assert tracer.updateGeometry("ground", ground, translation=(0, 0, 0), rotation=(0, 0, 0)) == 0
            assert tracer.updateGeometry("ben", ben, translation=lin, rotation=ang) == 0
            assert tracer.commitScene() == 0
            assert tracer.traceScene(k) == 0
            h, pts = tracer.getTraceCloud()
            A = lso.affine_from_components(lin, ang)
            ref = lso.scan(s, [(meshes["ground_v"], meshes["ground_t"], IDENT), (meshes["ben_v"], meshes["ben_t"], A)], ring_mode=1)
            assert h.width == ref["points"].shape[0] and h.seq == k
            assert np.array_equal(pts, ref["points"])
        tracer.close()
