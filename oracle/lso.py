"""ctypes front-end of the CPU ORACLE (oracle/lsb_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import
this module.  Nothing under lidarshooter_b200/ imports it.

It restates, on the CPU, the reference's per-frame trace path (SRC = ros_ws/src/lidarshooter/src):
updateGeometry (SRC/EmbreeTracer.cpp:262-288 -> SRC/MeshTransformer.cpp:142-205), commitScene
(SRC/EmbreeTracer.cpp:290-295), traceScene (SRC/EmbreeTracer.cpp:297-367) and the 32-byte point record
(SRC/XYZIRBytes.cpp:24-40).  See the header of lsb_oracle.c for the parity status ("parity unpinned"
per ray against Embree itself; point COUNTS 1668 / 1781 / 0 are pinned by the reference's tests).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liblsb_oracle.so")


def build(force=False):
    """Compile the oracle with the committed Makefile (gcc, -ffp-contract=off)."""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(
            os.path.join(_HERE, "lsb_oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        f32p, u32p, i32p, u8p = (C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_int32),
                                 C.POINTER(C.c_uint8))
        L.lso_quat_to_rotation.argtypes = [f32p, f32p]
        L.lso_inverse3.argtypes = [f32p, f32p]
        L.lso_affine_from_components.argtypes = [f32p, f32p, f32p]
        L.lso_transform_vertices.argtypes = [u8p, C.c_int64, C.c_int, f32p, f32p, f32p, C.c_int, f32p]
        L.lso_azimuth_step.argtypes = [C.c_float, C.c_float, C.c_uint]
        L.lso_azimuth_step.restype = C.c_float
        L.lso_ray_tables.argtypes = [f32p, C.c_int, C.c_float, C.c_float, C.c_int, f32p, f32p, f32p, f32p]
        L.lso_ray_dirs.argtypes = [f32p, f32p, C.c_int, f32p, f32p, C.c_int, f32p]
        L.lso_rotate_dirs.argtypes = [f32p, f32p, C.c_int64, f32p]
        L.lso_intersect_one.argtypes = [f32p, f32p, f32p, f32p, f32p, f32p]
        L.lso_intersect_one.restype = C.c_int
        L.lso_trace_brute.argtypes = [f32p, u32p, C.c_int64, f32p, C.c_int, f32p, C.c_int64, f32p, i32p]
        L.lso_bvh_build.argtypes = [f32p, u32p, C.c_int64]
        L.lso_bvh_build.restype = C.c_void_p
        L.lso_bvh_free.argtypes = [C.c_void_p]
        L.lso_bvh_node_count.argtypes = [C.c_void_p]
        L.lso_bvh_node_count.restype = C.c_int64
        L.lso_trace_bvh.argtypes = [C.c_void_p, f32p, C.c_int, f32p, C.c_int64, f32p, i32p]
        L.lso_trace_bvh_packet.argtypes = [C.c_void_p, f32p, C.c_int, f32p, C.c_int64, f32p, i32p]
        L.lso_trace_bvh_packet.restype = C.c_int
        L.lso_trace_bvh_second.argtypes = [C.c_void_p, f32p, C.c_int, f32p, C.c_int64, f32p, i32p, f32p, i32p]
        f64p = C.POINTER(C.c_double)
        L.lso_trace_bvh_f64.argtypes = [C.c_void_p, f32p, u32p, f32p, C.c_int, f32p, C.c_int64, f64p, i32p]
        L.lso_trace_brute_f64.argtypes = [f32p, u32p, C.c_int64, f32p, C.c_int, f32p, C.c_int64, f64p, i32p]
        L.lso_bary_f64.argtypes = [f32p, u32p, f32p, C.c_int, f32p, i32p, C.c_int64, f64p]
        L.lso_set_watertight.argtypes = [C.c_int]
        L.lso_get_watertight.restype = C.c_int
        L.lso_intensity.argtypes = [f32p, u32p, f32p, C.c_int64, C.c_int, C.c_float, f32p, i32p, f32p]
        L.lso_set_threads.argtypes = [C.c_int]
        L.lso_max_threads.restype = C.c_int
        L.lso_pack_points.argtypes = [f32p, f32p, i32p, C.c_int64, C.c_int, C.c_int, u8p]
        L.lso_pack_points.restype = C.c_int64
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


# ---------------------------------------------------------------------------------------------
# small pieces
# ---------------------------------------------------------------------------------------------
def quat_to_rotation(q_wxyz):
    q = _f32(q_wxyz)
    R = np.empty(9, np.float32)
    lib().lso_quat_to_rotation(_p(q, C.c_float), _p(R, C.c_float))
    return R


def inverse3(M):
    M = _f32(M).reshape(9)
    Mi = np.empty(9, np.float32)
    lib().lso_inverse3(_p(M, C.c_float), _p(Mi, C.c_float))
    return Mi


def affine_from_components(translation, euler_xyz):
    t, r = _f32(translation), _f32(euler_xyz)
    A = np.empty(12, np.float32)
    lib().lso_affine_from_components(_p(t, C.c_float), _p(r, C.c_float), _p(A, C.c_float))
    return A


IDENTITY_AFFINE = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)


def transform_vertices(verts, A, Rinv, T, mode=0, stride=None):
    """verts: (n,3) float32, or a uint8 blob with `stride` bytes per vertex (xyz f32 at offset 0)."""
    if stride is None:
        v = _f32(verts).reshape(-1, 3)
        n, stride, blob = v.shape[0], 12, v.view(np.uint8).reshape(-1)
    else:
        blob = np.ascontiguousarray(verts, dtype=np.uint8).reshape(-1)
        n = blob.size // stride
    A, Rinv, T = _f32(A).reshape(12), _f32(Rinv).reshape(9), _f32(T).reshape(3)
    out = np.empty((n, 3), np.float32)
    lib().lso_transform_vertices(_p(blob, C.c_uint8), n, stride, _p(A, C.c_float), _p(Rinv, C.c_float),
                                 _p(T, C.c_float), mode, _p(out, C.c_float))
    return out


def azimuth_step(begin, end, count):
    return np.float32(lib().lso_azimuth_step(float(begin), float(end), int(count)))


def ray_tables(elev_deg, az_begin, az_end, H):
    e = _f32(elev_deg)
    V = e.size
    step = azimuth_step(az_begin, az_end, H)
    st, ct = np.empty(V, np.float32), np.empty(V, np.float32)
    sp, cp = np.empty(H, np.float32), np.empty(H, np.float32)
    lib().lso_ray_tables(_p(e, C.c_float), V, float(np.float32(az_begin)), float(step), H, _p(st, C.c_float),
                         _p(ct, C.c_float), _p(sp, C.c_float), _p(cp, C.c_float))
    return st, ct, sp, cp


def ray_dirs(tables):
    st, ct, sp, cp = tables
    V, H = st.size, sp.size
    d = np.empty((V * H, 3), np.float32)
    lib().lso_ray_dirs(_p(st, C.c_float), _p(ct, C.c_float), V, _p(sp, C.c_float), _p(cp, C.c_float), H,
                       _p(d, C.c_float))
    return d


def rotate_dirs(R, dirs):
    R, dirs = _f32(R).reshape(9), _f32(dirs).reshape(-1, 3)
    out = np.empty_like(dirs)
    lib().lso_rotate_dirs(_p(R, C.c_float), _p(dirs, C.c_float), dirs.shape[0], _p(out, C.c_float))
    return out


def intersect_one(a, b, c, org, dirv):
    t = C.c_float(0)
    hit = lib().lso_intersect_one(_p(_f32(a), C.c_float), _p(_f32(b), C.c_float), _p(_f32(c), C.c_float),
                                  _p(_f32(org), C.c_float), _p(_f32(dirv), C.c_float), C.byref(t))
    return bool(hit), np.float32(t.value)


def _origins(orgs, nR):
    o = _f32(orgs)
    if o.size == 3:
        return o.reshape(3), 0
    o = o.reshape(-1, 3)
    assert o.shape[0] == nR
    return o, 3


def trace_brute(verts, tris, orgs, dirs):
    verts, dirs = _f32(verts).reshape(-1, 3), _f32(dirs).reshape(-1, 3)
    tris = np.ascontiguousarray(tris, dtype=np.uint32).reshape(-1, 3)
    nR = dirs.shape[0]
    o, stride = _origins(orgs, nR)
    t, tri = np.empty(nR, np.float32), np.empty(nR, np.int32)
    lib().lso_trace_brute(_p(verts, C.c_float), _p(tris, C.c_uint32), tris.shape[0], _p(o, C.c_float), stride,
                          _p(dirs, C.c_float), nR, _p(t, C.c_float), _p(tri, C.c_int32))
    return t, tri


def trace_brute_f64(verts, tris, orgs, dirs):
    verts, dirs = _f32(verts).reshape(-1, 3), _f32(dirs).reshape(-1, 3)
    tris = np.ascontiguousarray(tris, dtype=np.uint32).reshape(-1, 3)
    nR = dirs.shape[0]
    o, stride = _origins(orgs, nR)
    t, tri = np.empty(nR, np.float64), np.empty(nR, np.int32)
    lib().lso_trace_brute_f64(_p(verts, C.c_float), _p(tris, C.c_uint32), tris.shape[0], _p(o, C.c_float), stride,
                              _p(dirs, C.c_float), nR, _p(t, C.c_double), _p(tri, C.c_int32))
    return t, tri


def bary_f64(verts, tris, orgs, dirs, tri_ids):
    """(u, v, t, det) in float64 of every ray against triangle tri_ids[r], hit or not (NaN rows for id < 0)."""
    verts, dirs = _f32(verts).reshape(-1, 3), _f32(dirs).reshape(-1, 3)
    tris = np.ascontiguousarray(tris, dtype=np.uint32).reshape(-1, 3)
    ids = np.ascontiguousarray(tri_ids, dtype=np.int32)
    nR = dirs.shape[0]
    o, stride = _origins(orgs, nR)
    out = np.empty((nR, 4), np.float64)
    lib().lso_bary_f64(_p(verts, C.c_float), _p(tris, C.c_uint32), _p(o, C.c_float), stride, _p(dirs, C.c_float),
                       _p(ids, C.c_int32), nR, _p(out, C.c_double))
    return out


class Bvh:
    """Binned-SAH CPU BVH: stand-in for rtcCommitScene + rtcIntersect (SRC/EmbreeTracer.cpp:290-295, 472-480)."""

    def __init__(self, verts, tris):
        self._v = _f32(verts).reshape(-1, 3)
        self._t = np.ascontiguousarray(tris, dtype=np.uint32).reshape(-1, 3)
        self._h = lib().lso_bvh_build(_p(self._v, C.c_float), _p(self._t, C.c_uint32), self._t.shape[0])

    def trace(self, orgs, dirs):
        dirs = _f32(dirs).reshape(-1, 3)
        nR = dirs.shape[0]
        o, stride = _origins(orgs, nR)
        t, tri = np.empty(nR, np.float32), np.empty(nR, np.int32)
        lib().lso_trace_bvh(self._h, _p(o, C.c_float), stride, _p(dirs, C.c_float), nR, _p(t, C.c_float),
                            _p(tri, C.c_int32))
        return t, tri

    def trace_second(self, orgs, dirs, t1, tri1):
        """Second return per ray: the smallest (t, triangle id) greater than the first hit (t1, tri1); (+inf, -1) if none."""
        dirs = _f32(dirs).reshape(-1, 3)
        nR = dirs.shape[0]
        o, stride = _origins(orgs, nR)
        t1 = _f32(t1); tri1 = np.ascontiguousarray(tri1, dtype=np.int32)
        t, tri = np.empty(nR, np.float32), np.empty(nR, np.int32)
        lib().lso_trace_bvh_second(self._h, _p(o, C.c_float), stride, _p(dirs, C.c_float), nR, _p(t1, C.c_float),
                                   _p(tri1, C.c_int32), _p(t, C.c_float), _p(tri, C.c_int32))
        return t, tri

    def trace_packet(self, orgs, dirs):
        """8-ray AVX2 packets through the same BVH (the CPU baseline closer to rtcIntersect16); bit-identical to trace().
        Returns None when the host has no AVX2 or the watertight mode is on."""
        dirs = _f32(dirs).reshape(-1, 3)
        nR = dirs.shape[0]
        o, stride = _origins(orgs, nR)
        t, tri = np.empty(nR, np.float32), np.empty(nR, np.int32)
        ok = lib().lso_trace_bvh_packet(self._h, _p(o, C.c_float), stride, _p(dirs, C.c_float), nR, _p(t, C.c_float),
                                        _p(tri, C.c_int32))
        return (t, tri) if ok else None

    def trace_f64(self, orgs, dirs):
        """Float64 second opinion (textbook Moeller-Trumbore in double on the original fp32 vertices)."""
        dirs = _f32(dirs).reshape(-1, 3)
        nR = dirs.shape[0]
        o, stride = _origins(orgs, nR)
        t, tri = np.empty(nR, np.float64), np.empty(nR, np.int32)
        lib().lso_trace_bvh_f64(self._h, _p(self._v, C.c_float), _p(self._t, C.c_uint32), _p(o, C.c_float), stride,
                                _p(dirs, C.c_float), nR, _p(t, C.c_double), _p(tri, C.c_int32))
        return t, tri

    def node_count(self):
        return int(lib().lso_bvh_node_count(self._h))

    def close(self):
        if self._h:
            lib().lso_bvh_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def pack_points(dirs_s, t, tri, H, ring_mode=0, intensity=None):
    dirs_s, t = _f32(dirs_s).reshape(-1, 3), _f32(t)
    tri = np.ascontiguousarray(tri, dtype=np.int32)
    out = np.zeros(32 * t.size, np.uint8)
    n = lib().lso_pack_points(_p(dirs_s, C.c_float), _p(t, C.c_float), _p(tri, C.c_int32), t.size, int(H),
                              int(ring_mode), _p(out, C.c_uint8))
    pts = out[:32 * n].reshape(n, 32)
    if intensity is not None:                 # per-ray intensities of an intensity model (default: the reference's 64.0)
        pts.view(np.float32).reshape(n, 8)[:, 4] = _f32(intensity)[tri >= 0]
    return pts


class watertight:
    """Context manager: the oracle's tracers use the watertight Woop-Benthin-Wald test inside (include/lsb200.h:
    LSB200_OPT_WATERTIGHT) instead of the reference's Embree-ordered Moeller-Trumbore."""

    def __init__(self, on=True):
        self.on = on

    def __enter__(self):
        self.prev = lib().lso_get_watertight()
        lib().lso_set_watertight(1 if self.on else 0)
        return self

    def __exit__(self, *a):
        lib().lso_set_watertight(self.prev)


def intensity(verts, tris, dirs, t, tri, model=1, min_intensity=0.0):
    """Intensity model / filter of include/lsb200.h.  Returns (t, tri, intensity) -- new arrays."""
    verts, dirs = _f32(verts).reshape(-1, 3), _f32(dirs).reshape(-1, 3)
    tris = np.ascontiguousarray(tris, dtype=np.uint32).reshape(-1, 3)
    t2, tri2 = _f32(t).copy(), np.ascontiguousarray(tri, dtype=np.int32).copy()
    out = np.zeros(t2.size, np.float32)
    lib().lso_intensity(_p(verts, C.c_float), _p(tris, C.c_uint32), _p(dirs, C.c_float), t2.size, int(model),
                        float(min_intensity), _p(t2, C.c_float), _p(tri2, C.c_int32), _p(out, C.c_float))
    return t2, tri2, out


def set_threads(n):
    lib().lso_set_threads(int(n))


def max_threads():
    return int(lib().lso_max_threads())


# ---------------------------------------------------------------------------------------------
# whole-path restatement: one scan of one device
# ---------------------------------------------------------------------------------------------
class Sensor:
    """The numbers LidarDevice holds for the path (SRC/LidarDevice.cpp:598-617, 788-816)."""

    def __init__(self, elev_deg, az_begin, az_end, H, q_wxyz=(1, 0, 0, 0), t_xyz=(0, 0, 0)):
        self.elev = _f32(elev_deg)
        self.V, self.H = int(self.elev.size), int(H)
        self.az_begin, self.az_end = np.float32(az_begin), np.float32(az_end)
        self.q = _f32(q_wxyz)
        self.T = _f32(t_xyz)
        self.R = quat_to_rotation(self.q)
        self.Rinv = inverse3(self.R)
        self.tables = ray_tables(self.elev, self.az_begin, self.az_end, self.H)

    @property
    def n_rays(self):
        return self.V * self.H


def concat_scene(sensor, meshes, mode=0):
    """meshes: list of (verts (n,3) f32, tris (m,3) u32, affine12) in ascending geometry-id order.
    Returns transformed vertices and triangles with global vertex numbering (global triangle id =
    position in the concatenation)."""
    vs, ts, off = [], [], 0
    for verts, tris, A in meshes:
        v = transform_vertices(verts, A, sensor.Rinv, sensor.T, mode)
        vs.append(v)
        ts.append(np.asarray(tris, np.uint32).reshape(-1, 3) + np.uint32(off))
        off += v.shape[0]
    if not vs:
        return np.zeros((0, 3), np.float32), np.zeros((0, 3), np.uint32)
    return np.concatenate(vs), np.concatenate(ts)


def azimuth_degrees(sensor):
    """Azimuth of every step as the ray generator computes it: begin + step * float(h) in float32 (SRC/LidarDevice.cpp:306)."""
    step = azimuth_step(sensor.az_begin, sensor.az_end, sensor.H)
    return (np.float32(sensor.az_begin) + step * np.arange(sensor.H, dtype=np.float32)).astype(np.float32)


def apply_limits(sensor, t, tri, max_range=0.0, az_window=(-180.0, 180.0)):
    """Sensor limits of include/lsb200.h: lsb200_set_sensor_limits (semantics defined there; the reference only carries
    the config keys).  Returns new (t, tri)."""
    t, tri = t.copy(), tri.copy()
    w = azimuth_degrees(sensor).copy()
    for _ in range(4):
        w = np.where(w > np.float32(180), w - np.float32(360), w).astype(np.float32)
        w = np.where(w <= np.float32(-180), w + np.float32(360), w).astype(np.float32)
    fired = (w >= np.float32(az_window[0])) & (w <= np.float32(az_window[1]))
    if az_window[0] <= -180.0 and az_window[1] >= 180.0:
        fired[:] = True
    dead = np.tile(~fired, sensor.V)
    if max_range and max_range > 0:
        dead |= ~(t <= np.float32(max_range))
    t[dead] = np.inf
    tri[dead] = -1
    return t, tri


def scan(sensor, meshes, mode=0, use_bvh=None, ring_mode=0, max_range=0.0, az_window=(-180.0, 180.0)):
    """One traceScene.  mode 0 = sensor frame (the reference's arithmetic: the SCENE is moved into the
    sensor frame, rays start at 0); mode 1 = world frame (rays are moved: o = T, d = R d_s).
    Returns dict(t, tri, points) with points in ray-id order."""
    verts, tris = concat_scene(sensor, meshes, mode)
    dirs_s = ray_dirs(sensor.tables)
    if mode == 0:
        org, dirs = np.zeros(3, np.float32), dirs_s
    else:
        org, dirs = sensor.T, rotate_dirs(sensor.R, dirs_s)
    if use_bvh is None:
        use_bvh = tris.shape[0] * dirs.shape[0] > 50_000_000
    if tris.shape[0] == 0:
        t = np.full(dirs.shape[0], np.inf, np.float32)
        tri = np.full(dirs.shape[0], -1, np.int32)
    elif use_bvh:
        b = Bvh(verts, tris)
        t, tri = b.trace(org, dirs)
        b.close()
    else:
        t, tri = trace_brute(verts, tris, org, dirs)
    if (max_range and max_range > 0) or az_window[0] > -180.0 or az_window[1] < 180.0:
        t, tri = apply_limits(sensor, t, tri, max_range, az_window)
    pts = pack_points(dirs_s, t, tri, sensor.H, ring_mode)
    return {"t": t, "tri": tri, "points": pts, "verts": verts, "tris": tris}
