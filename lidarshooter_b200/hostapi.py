"""ctypes binding of include/lsb200_host.h (liblsb200_host.so): the C++ host layer -- LidarDevice, mesh readers,
B200Tracer (the ITracer drop-in), AffineMesh / Trajectory -- as the tests drive it."""
import ctypes as C
import os

import numpy as np

from . import capi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liblsb200_host.so")

SYMBOLS = [
    "lsbh_device_create", "lsbh_device_create_synthetic", "lsbh_device_destroy", "lsbh_device_sensor_uid",
    "lsbh_device_total_rays", "lsbh_device_total_channels", "lsbh_device_scan_ray_count", "lsbh_device_transform",
    "lsbh_device_channels", "lsbh_device_next_ray", "lsbh_device_reset", "lsbh_device_all_rays",
    "lsbh_device_origin_to_sensor", "lsbh_device_init_message", "lsbh_mesh_load", "lsbh_mesh_from_arrays",
    "lsbh_mesh_destroy", "lsbh_mesh_counts", "lsbh_mesh_copy", "lsbh_tracer_create", "lsbh_tracer_destroy",
    "lsbh_tracer_add_geometry", "lsbh_tracer_remove_geometry", "lsbh_tracer_geometry_count", "lsbh_tracer_geometry_id",
    "lsbh_tracer_update_geometry", "lsbh_tracer_update_geometry_euler", "lsbh_tracer_commit_scene",
    "lsbh_tracer_trace_scene", "lsbh_tracer_set_flags", "lsbh_tracer_cloud", "lsbh_tracer_trace_raw",
    "lsbh_tracer_serialize_cloud", "lsbh_cloud_reserialize", "lsbh_tracer_append_cloud_to_file",
    "lsbh_device_limits", "lsbh_tracer_set_sensor_limits", "lsbh_tracer_set_intensity", "lsbh_tracer_set_watertight",
    "lsbh_tracer_log_count", "lsbh_tracer_log_line",
    "lsbh_trajectory_poses",
]

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    capi.lib()      # liblsb200.so first (the host library links against it)
    if not os.path.exists(LIB_PATH):
        raise OSError(f"{LIB_PATH} not found: build it with `make -C lidarshooter_b200`")
    L = C.CDLL(LIB_PATH)
    vp, f32p, u32p, i32p, cp = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_int32), C.c_char_p
    sig = {
        "lsbh_device_create": (vp, [cp, cp, cp, C.c_size_t]),
        "lsbh_device_create_synthetic": (vp, [f32p, C.c_int, C.c_float, C.c_float, C.c_uint, f32p, f32p]),
        "lsbh_device_destroy": (None, [vp]),
        "lsbh_device_sensor_uid": (cp, [vp]),
        "lsbh_device_total_rays": (C.c_uint, [vp]),
        "lsbh_device_total_channels": (C.c_uint, [vp]),
        "lsbh_device_scan_ray_count": (C.c_uint, [vp]),
        "lsbh_device_transform": (C.c_int, [vp, f32p, f32p, f32p, f32p]),
        "lsbh_device_channels": (C.c_int, [vp, f32p, C.c_int, f32p, f32p, f32p]),
        "lsbh_device_next_ray": (C.c_int, [vp, f32p, i32p, i32p]),
        "lsbh_device_reset": (None, [vp]),
        "lsbh_device_all_rays": (C.c_int, [vp, f32p, C.c_size_t]),
        "lsbh_device_origin_to_sensor": (None, [vp, f32p, C.c_int]),
        "lsbh_device_init_message": (C.c_int, [vp, C.c_int, C.POINTER(capi.CloudHeader), cp, C.c_size_t, cp, C.c_size_t]),
        "lsbh_mesh_load": (vp, [cp]),
        "lsbh_mesh_from_arrays": (vp, [f32p, C.c_size_t, u32p, C.c_size_t, C.c_int]),
        "lsbh_mesh_destroy": (None, [vp]),
        "lsbh_mesh_counts": (C.c_int, [vp, i32p, i32p, i32p]),
        "lsbh_mesh_copy": (C.c_int, [vp, f32p, u32p, C.c_int]),
        "lsbh_tracer_create": (vp, [vp, C.c_int, cp, C.c_size_t]),
        "lsbh_tracer_destroy": (None, [vp]),
        "lsbh_tracer_add_geometry": (C.c_int, [vp, cp, C.c_int, C.c_int, C.c_int]),
        "lsbh_tracer_remove_geometry": (C.c_int, [vp, cp]),
        "lsbh_tracer_geometry_count": (C.c_long, [vp]),
        "lsbh_tracer_geometry_id": (C.c_int, [vp, cp]),
        "lsbh_tracer_update_geometry": (C.c_int, [vp, cp, vp, f32p]),
        "lsbh_tracer_update_geometry_euler": (C.c_int, [vp, cp, vp, f32p, f32p]),
        "lsbh_tracer_commit_scene": (C.c_int, [vp]),
        "lsbh_tracer_trace_scene": (C.c_int, [vp, C.c_uint32]),
        "lsbh_tracer_set_flags": (C.c_int, [vp, C.c_int, C.c_int, C.c_int]),
        "lsbh_tracer_cloud": (C.c_int, [vp, C.POINTER(capi.CloudHeader), C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_size_t)]),
        "lsbh_tracer_trace_raw": (C.c_int, [vp, f32p, i32p]),
        "lsbh_trajectory_poses": (C.c_int, [cp, C.c_float, f32p, C.c_int]),
        "lsbh_tracer_serialize_cloud": (C.c_long, [vp, C.c_double, C.c_int, C.POINTER(C.c_uint8), C.c_size_t]),
        "lsbh_cloud_reserialize": (C.c_long, [C.POINTER(C.c_uint8), C.c_size_t, C.POINTER(C.c_uint8), C.c_size_t,
                                              C.POINTER(capi.CloudHeader), C.POINTER(C.c_double)]),
        "lsbh_tracer_append_cloud_to_file": (C.c_int, [vp, cp, C.c_int]),
        "lsbh_device_limits": (C.c_int, [vp, f32p, f32p, f32p]),
        "lsbh_tracer_set_sensor_limits": (C.c_int, [vp, C.c_int]),
        "lsbh_tracer_set_intensity": (C.c_int, [vp, C.c_int, C.c_int]),
        "lsbh_tracer_set_watertight": (C.c_int, [vp, C.c_int]),
        "lsbh_tracer_log_count": (C.c_int, [vp]),
        "lsbh_tracer_log_line": (C.c_int, [vp, C.c_int, C.POINTER(C.c_int), C.c_char_p, C.c_size_t]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class LidarDevice:
    def __init__(self, config=None, sensor_uid=None, synthetic=None):
        L = lib()
        self._L = L
        if synthetic is not None:
            elev, b, e, H, q, t = synthetic
            elev = np.ascontiguousarray(elev, np.float32)
            q, t = np.ascontiguousarray(q, np.float32), np.ascontiguousarray(t, np.float32)
            self._h = L.lsbh_device_create_synthetic(_fp(elev), elev.size, float(b), float(e), int(H), _fp(q), _fp(t))
        else:
            err = C.create_string_buffer(512)
            self._h = L.lsbh_device_create(config.encode(), (sensor_uid or "").encode(), err, 512)
            if not self._h:
                raise RuntimeError(err.value.decode())

    def close(self):
        if self._h:
            self._L.lsbh_device_destroy(self._h)
            self._h = None

    def limits(self):
        """(detection_radius, az_lo, az_hi) of the device config."""
        r, lo, hi = C.c_float(), C.c_float(), C.c_float()
        self._L.lsbh_device_limits(self._h, C.byref(r), C.byref(lo), C.byref(hi))
        return r.value, lo.value, hi.value

    sensor_uid = property(lambda s: s._L.lsbh_device_sensor_uid(s._h).decode())
    total_rays = property(lambda s: s._L.lsbh_device_total_rays(s._h))
    total_channels = property(lambda s: s._L.lsbh_device_total_channels(s._h))
    scan_ray_count = property(lambda s: s._L.lsbh_device_scan_ray_count(s._h))

    def transform(self):
        q, t, R, Ri = np.empty(4, np.float32), np.empty(3, np.float32), np.empty(9, np.float32), np.empty(9, np.float32)
        self._L.lsbh_device_transform(self._h, _fp(q), _fp(t), _fp(R), _fp(Ri))
        return q, t, R, Ri

    def channels(self):
        v = np.empty(self.total_channels, np.float32)
        b, e, s = C.c_float(), C.c_float(), C.c_float()
        self._L.lsbh_device_channels(self._h, _fp(v), v.size, C.byref(b), C.byref(e), C.byref(s))
        return v, np.float32(b.value), np.float32(e.value), np.float32(s.value)

    def next_ray(self):
        d = np.empty(3, np.float32)
        vi, hi = C.c_int32(), C.c_int32()
        state = self._L.lsbh_device_next_ray(self._h, _fp(d), C.byref(vi), C.byref(hi))
        return state, d, vi.value, hi.value

    def reset(self):
        self._L.lsbh_device_reset(self._h)

    def all_rays(self):
        d = np.empty((self.total_rays, 3), np.float32)
        assert self._L.lsbh_device_all_rays(self._h, _fp(d), d.shape[0]) == d.shape[0]
        return d

    def origin_to_sensor(self, p, inverse=False):
        q = np.ascontiguousarray(p, np.float32).copy()
        self._L.lsbh_device_origin_to_sensor(self._h, _fp(q), 1 if inverse else 0)
        return q

    def init_message(self, frame):
        h = capi.CloudHeader()
        fid, fields = C.create_string_buffer(128), C.create_string_buffer(512)
        self._L.lsbh_device_init_message(self._h, frame, C.byref(h), fid, 128, fields, 512)
        ft = [tuple(x.split(":")) for x in fields.value.decode().split(";") if x]
        return h, fid.value.decode(), [(n, int(o), int(d), int(c)) for n, o, d, c in ft]


class Mesh:
    def __init__(self, path=None, arrays=None):
        L = lib()
        self._L = L
        if arrays is not None:
            v, e = arrays
            v = np.ascontiguousarray(v, np.float32).reshape(-1, 3)
            e = np.ascontiguousarray(e, np.uint32)
            self._h = L.lsbh_mesh_from_arrays(_fp(v), v.shape[0], e.ctypes.data_as(C.POINTER(C.c_uint32)), e.shape[0], e.shape[1])
        else:
            self._h = L.lsbh_mesh_load(path.encode())
            if not self._h:
                raise RuntimeError(f"cannot load {path}")

    def counts(self):
        a, b, c = C.c_int32(), C.c_int32(), C.c_int32()
        self._L.lsbh_mesh_counts(self._h, C.byref(a), C.byref(b), C.byref(c))
        return a.value, b.value, c.value

    def arrays(self, verts_per_elem=3):
        nv, npoly, _ = self.counts()
        v, e = np.empty((nv, 3), np.float32), np.empty((npoly, verts_per_elem), np.uint32)
        assert self._L.lsbh_mesh_copy(self._h, _fp(v), e.ctypes.data_as(C.POINTER(C.c_uint32)), verts_per_elem) == 0
        return v, e

    def close(self):
        if self._h:
            self._L.lsbh_mesh_destroy(self._h)
            self._h = None


class B200Tracer:
    """lidarshooter::B200Tracer through the C flattening; method names follow ITracer (SRC/ITracer.hpp:50-94)."""

    def __init__(self, device, cuda_device=0):
        L = lib()
        self._L, self._dev = L, device
        err = C.create_string_buffer(512)
        self._h = L.lsbh_tracer_create(device._h, cuda_device, err, 512)
        if not self._h:
            raise RuntimeError(err.value.decode())

    def close(self):
        if self._h:
            self._L.lsbh_tracer_destroy(self._h)
            self._h = None

    def addGeometry(self, name, gtype, nv, ne):
        return self._L.lsbh_tracer_add_geometry(self._h, name.encode(), gtype, nv, ne)

    def removeGeometry(self, name):
        return self._L.lsbh_tracer_remove_geometry(self._h, name.encode())

    def getGeometryCount(self):
        return self._L.lsbh_tracer_geometry_count(self._h)

    def getGeometryId(self, name):
        return self._L.lsbh_tracer_geometry_id(self._h, name.encode())

    def updateGeometry(self, name, mesh, affine=None, translation=None, rotation=None):
        mh = mesh._h if mesh is not None else None
        if translation is not None:
            t, r = np.ascontiguousarray(translation, np.float32), np.ascontiguousarray(rotation, np.float32)
            return self._L.lsbh_tracer_update_geometry_euler(self._h, name.encode(), mh, _fp(t), _fp(r))
        A = np.ascontiguousarray(affine if affine is not None else [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)
        return self._L.lsbh_tracer_update_geometry(self._h, name.encode(), mh, _fp(A))

    def commitScene(self):
        return self._L.lsbh_tracer_commit_scene(self._h)

    def traceScene(self, frame):
        return self._L.lsbh_tracer_trace_scene(self._h, frame)

    def setFlags(self, ordered=False, ring_zero=True, world_points=False):
        """ring_zero=True is the tracer's default (the Embree path's bytes, SRC/EmbreeTracer.cpp:333-334); False = ring = channel."""
        self._L.lsbh_tracer_set_flags(self._h, int(ordered), int(ring_zero), int(world_points))

    def setIntensity(self, model=0, apply_config_filter=False):
        return self._L.lsbh_tracer_set_intensity(self._h, int(model), int(apply_config_filter))

    def setWatertight(self, on=True):
        return self._L.lsbh_tracer_set_watertight(self._h, int(on))

    def logLines(self):
        """(level, text) pairs the tracer wrote to the spdlog logger it was created with."""
        out = []
        for i in range(self._L.lsbh_tracer_log_count(self._h)):
            lv, buf = C.c_int(0), C.create_string_buffer(600)
            if self._L.lsbh_tracer_log_line(self._h, i, C.byref(lv), buf, 600) == 0:
                out.append((lv.value, buf.value.decode(errors="replace")))
        return out

    def getTraceCloud(self):
        h = capi.CloudHeader()
        data, n = C.POINTER(C.c_uint8)(), C.c_size_t()
        self._L.lsbh_tracer_cloud(self._h, C.byref(h), C.byref(data), C.byref(n))
        buf = np.ctypeslib.as_array(data, shape=(n.value,)).copy() if n.value else np.zeros(0, np.uint8)
        return h, buf.reshape(-1, 32)

    def getTraceCloudWidth(self):
        h = capi.CloudHeader()
        data, n = C.POINTER(C.c_uint8)(), C.c_size_t()
        self._L.lsbh_tracer_cloud(self._h, C.byref(h), C.byref(data), C.byref(n))
        return int(h.width)

    def setSensorLimits(self, on=True):
        return self._L.lsbh_tracer_set_sensor_limits(self._h, int(on))

    def serializeCloud(self, stamp=-1.0, fix_row_step=False):
        """ROS-1 wire bytes of the traced sensor_msgs/PointCloud2 (host/CloudSerializer.hpp)."""
        n = self._L.lsbh_tracer_serialize_cloud(self._h, float(stamp), int(fix_row_step), None, 0)
        if n < 0:
            raise RuntimeError("serialize failed")
        out = np.empty(n, np.uint8)
        m = self._L.lsbh_tracer_serialize_cloud(self._h, float(stamp), int(fix_row_step), out.ctypes.data_as(C.POINTER(C.c_uint8)), n)
        assert m == n
        return out

    def appendCloudToFile(self, path, fix_row_step=False):
        return self._L.lsbh_tracer_append_cloud_to_file(self._h, str(path).encode(), int(fix_row_step))

    def traceRaw(self):
        R = self._dev.total_rays
        t, tri = np.empty(R, np.float32), np.empty(R, np.int32)
        rc = self._L.lsbh_tracer_trace_raw(self._h, _fp(t), tri.ctypes.data_as(C.POINTER(C.c_int32)))
        return rc, t, tri


def reserialize_cloud(msg):
    """Parses ROS-1 PointCloud2 wire bytes with the C++ reader and serialises them again.
    Returns (bytes, header, stamp); raises on malformed input."""
    L = lib()
    msg = np.ascontiguousarray(msg, np.uint8)
    h, stamp = capi.CloudHeader(), C.c_double()
    out = np.empty(max(msg.size, 1), np.uint8)
    n = L.lsbh_cloud_reserialize(msg.ctypes.data_as(C.POINTER(C.c_uint8)), msg.size, out.ctypes.data_as(C.POINTER(C.c_uint8)),
                                 out.size, C.byref(h), C.byref(stamp))
    if n < 0:
        raise ValueError("malformed PointCloud2 message")
    return out[:n], h, stamp.value


def trajectory_poses(path, period=0.1):
    L = lib()
    n = L.lsbh_trajectory_poses(path.encode(), float(period), None, 0)
    if n < 0:
        raise RuntimeError(f"cannot read {path}")
    out = np.empty((n, 6), np.float32)
    L.lsbh_trajectory_poses(path.encode(), float(period), _fp(out), n)
    return out
