"""ctypes binding of the C ABI in include/lsb200.h (liblsb200.so).

This is plumbing for tests and bench.py: every call goes straight into the CUDA library.  There is no CPU
fallback: if liblsb200.so is missing or no CUDA device is present the import / lsb200_create fails loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LSB200_LIB") or os.path.join(_HERE, "liblsb200.so")   # override: A/B builds of tools/
# the same library built with -DLSB_LAB: plus every measured-and-rejected variant (A/B tests and tools/ only)
LAB_LIB_PATH = os.environ.get("LSB200_LAB_LIB") or os.path.join(_HERE, "liblsb200_lab.so")

GEOMETRY_TRIANGLE, GEOMETRY_QUAD = 0, 1
FRAME_SENSOR, FRAME_WORLD = 0, 1
COMMIT_AUTO, COMMIT_REBUILD, COMMIT_REFIT = 0, 1, 2
TRACE_ORDERED, TRACE_RING_ZERO, TRACE_XYZI16, TRACE_NO_POINTS, TRACE_WORLD_POINTS, TRACE_RANGE_IMAGE = 1, 2, 4, 8, 16, 32
OPT_TRAVERSAL, OPT_LEAF_SIZE, OPT_TOP_NODES, OPT_STAGE_TIMING, OPT_CTAS_PER_SM = 1, 2, 3, 4, 5
OPT_TRACE_STATS, OPT_MORTON_UNIFORM, OPT_ORIGIN_CHAIN, OPT_TILE_WIDTH, OPT_SMALL_BUILD, OPT_MORTON_LAYOUT, OPT_MORTON_ASPECT, OPT_FIT_RMQ = 6, 7, 8, 9, 10, 11, 12, 13
OPT_TILE_ENTRIES, OPT_TILE_VISITS, OPT_TILE_MIN_UNITS, OPT_SUPER_TILES, OPT_WATERTIGHT, OPT_INTENSITY_MODEL, OPT_FUSED_TREE, OPT_LANE_REFILL, OPT_RETURNS, OPT_SMALL_TRACE_UNITS = 14, 15, 16, 17, 18, 19, 20, 21, 22, 23
LOG_TRACE, LOG_DEBUG, LOG_INFO, LOG_WARN, LOG_ERROR = 0, 1, 2, 3, 4
SHARD_FRAMES, SHARD_SENSORS = 0, 1
SORT_STATIC_MAX_PAIRS = 1 << 22          # LSB200_SORT_STATIC_MAX_PAIRS
ERR_INVALID, ERR_EMPTY_SCENE, ERR_CUDA, ERR_CAPACITY, ERR_STATE, ERR_NOMEM = -1, -1, -2, -3, -4, -5

# every symbol include/lsb200.h declares (tests/test_abi.py checks the library exports each one)
SYMBOLS = [
    "lsb200_create", "lsb200_destroy", "lsb200_set_stream", "lsb200_set_frame_mode", "lsb200_set_option",
    "lsb200_last_error", "lsb200_version", "lsb200_set_sensor", "lsb200_set_sensor_limits", "lsb200_get_ray_count", "lsb200_add_geometry",
    "lsb200_remove_geometry", "lsb200_get_geometry_count", "lsb200_get_geometry_id", "lsb200_update_geometry",
    "lsb200_update_geometry_euler", "lsb200_update_transform", "lsb200_affine_from_components",
    "lsb200_commit_scene", "lsb200_trace_scene", "lsb200_trace_device", "lsb200_trace_batch", "lsb200_trace_raw",
    "lsb200_device_results", "lsb200_fetch_points", "lsb200_fetch_batch", "lsb200_synchronize", "lsb200_run_frames", "lsb200_init_message", "lsb200_get_stats",
    "lsb200_reset_counters", "lsb200_get_trace_counters", "lsb200_debug_copy_bvh", "lsb200_debug_sort_pairs", "lsb200_debug_copy_vertices",
    "lsb200_set_log_callback", "lsb200_set_sensor_intensity_filter", "lsb200_expand_range_image", "lsb200_fetch_intensity",
    "lsb200_set_sensor_pose", "lsb200_alloc_host", "lsb200_free_host", "lsb200_device_numa_node", "lsb200_clone_scene", "lsb200_run_batch", "lsb200_fetch_second_returns", "lsb200_sort_pairs_static",
]

LOG_FN = C.CFUNCTYPE(None, C.c_int, C.c_char_p, C.c_void_p)


class CloudHeader(C.Structure):
    _fields_ = [("seq", C.c_uint32), ("height", C.c_uint32), ("width", C.c_uint32), ("point_step", C.c_uint32),
                ("row_step", C.c_uint32), ("is_bigendian", C.c_uint8), ("is_dense", C.c_uint8),
                ("n_fields", C.c_uint8), ("reserved", C.c_uint8)]


class Frame(C.Structure):
    _fields_ = [("affines", C.POINTER(C.c_float)), ("verts", C.POINTER(C.c_void_p)), ("out_points", C.POINTER(C.c_void_p)),
                ("capacities", C.POINTER(C.c_size_t)), ("n_pts", C.POINTER(C.c_uint32))]


class Stats(C.Structure):
    _fields_ = [("n_vertices", C.c_int64), ("n_triangles", C.c_int64), ("n_nodes", C.c_int64), ("n_leaves", C.c_int64),
                ("last_commit_was_rebuild", C.c_int32), ("top_nodes", C.c_int32), ("trace_grid", C.c_int32),
                ("trace_block", C.c_int32),
                ("ms_transform", C.c_float), ("ms_morton", C.c_float), ("ms_sort", C.c_float),
                ("ms_hierarchy", C.c_float), ("ms_fit", C.c_float), ("ms_toplevel", C.c_float),
                ("ms_trace", C.c_float), ("ms_compact", C.c_float),
                ("kernel_launches", C.c_int64), ("scene_lo", C.c_float * 3), ("scene_hi", C.c_float * 3),
                ("ms_tile_lists", C.c_float), ("reserved0", C.c_int32)]


_libs = {}


def lib(lab=False):
    """Loads liblsb200.so -- or, lab=True, liblsb200_lab.so -- (raises OSError with the build hint when it is absent)."""
    if lab in _libs:
        return _libs[lab]
    path = LAB_LIB_PATH if lab else LIB_PATH
    if not os.path.exists(path):
        raise OSError(f"{path} not found: build it with `make -C lidarshooter_b200` "
                      "(or __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(path)
    vp, f32p, u32p, i32p = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_int32)
    L.lsb200_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.lsb200_destroy.argtypes = [vp]
    L.lsb200_set_stream.argtypes = [vp, vp]
    L.lsb200_set_frame_mode.argtypes = [vp, C.c_int]
    L.lsb200_set_option.argtypes = [vp, C.c_int, C.c_int]
    L.lsb200_last_error.argtypes = [vp]
    L.lsb200_last_error.restype = C.c_char_p
    L.lsb200_version.restype = C.c_char_p
    L.lsb200_set_sensor.argtypes = [vp, C.c_int, f32p, C.c_int, C.c_float, C.c_float, C.c_int, f32p, f32p]
    L.lsb200_set_sensor_limits.argtypes = [vp, C.c_int, C.c_float, C.c_float, C.c_float]
    L.lsb200_get_ray_count.argtypes = [vp, C.c_int]
    L.lsb200_add_geometry.argtypes = [vp, C.c_char_p, C.c_int, C.c_int, C.c_int]
    L.lsb200_remove_geometry.argtypes = [vp, C.c_char_p]
    L.lsb200_get_geometry_count.argtypes = [vp]
    L.lsb200_get_geometry_id.argtypes = [vp, C.c_char_p]
    L.lsb200_update_geometry.argtypes = [vp, C.c_char_p, vp, C.c_int, vp, f32p]
    L.lsb200_update_geometry_euler.argtypes = [vp, C.c_char_p, vp, C.c_int, vp, f32p, f32p]
    L.lsb200_update_transform.argtypes = [vp, C.c_char_p, f32p]
    L.lsb200_affine_from_components.argtypes = [f32p, f32p, f32p]
    L.lsb200_affine_from_components.restype = None
    L.lsb200_commit_scene.argtypes = [vp, C.c_int]
    L.lsb200_trace_scene.argtypes = [vp, C.c_int, C.c_uint32, vp, C.c_size_t, u32p, C.c_uint32]
    L.lsb200_trace_device.argtypes = [vp, C.c_int, C.c_uint32]
    L.lsb200_trace_batch.argtypes = [vp, C.c_int, C.c_int, C.c_uint32]
    L.lsb200_trace_raw.argtypes = [vp, C.c_int, f32p, i32p]
    L.lsb200_fetch_second_returns.argtypes = [vp, C.c_int, f32p, i32p]
    L.lsb200_device_results.argtypes = [vp, C.c_int, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    L.lsb200_fetch_points.argtypes = [vp, C.c_int, vp, C.c_size_t, u32p, C.c_uint32]
    L.lsb200_fetch_batch.argtypes = [vp, C.c_int, C.c_int, C.POINTER(vp), C.POINTER(C.c_size_t), u32p, C.c_uint32]
    L.lsb200_synchronize.argtypes = [vp]
    L.lsb200_run_frames.argtypes = [vp, C.POINTER(Frame), C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint32]
    L.lsb200_init_message.argtypes = [vp, C.c_uint32, C.POINTER(CloudHeader)]
    L.lsb200_get_stats.argtypes = [vp, C.POINTER(Stats)]
    L.lsb200_reset_counters.argtypes = [vp]
    L.lsb200_get_trace_counters.argtypes = [vp, C.POINTER(C.c_uint64), C.c_int]
    L.lsb200_debug_copy_bvh.argtypes = [vp, f32p, C.c_int64, i32p, C.c_int64, i32p]
    L.lsb200_debug_sort_pairs.argtypes = [vp, u32p, u32p, C.c_int64]
    L.lsb200_debug_copy_vertices.argtypes = [vp, f32p, C.c_int64]
    L.lsb200_set_log_callback.argtypes = [vp, LOG_FN, vp]
    L.lsb200_set_sensor_intensity_filter.argtypes = [vp, C.c_int, C.c_float]
    L.lsb200_set_sensor_pose.argtypes = [vp, C.c_int, f32p, f32p]
    L.lsb200_expand_range_image.argtypes = [vp, C.c_int, f32p, vp, C.c_size_t, u32p, C.c_uint32]
    L.lsb200_fetch_intensity.argtypes = [vp, C.c_int, f32p]
    L.lsb200_alloc_host.argtypes = [vp, C.c_size_t, C.POINTER(vp)]
    L.lsb200_free_host.argtypes = [vp, vp]
    L.lsb200_device_numa_node.argtypes = [vp]
    L.lsb200_clone_scene.argtypes = [vp, C.c_int, C.POINTER(vp)]
    L.lsb200_run_batch.argtypes = [C.POINTER(vp), C.c_int, C.POINTER(Frame), C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_int]
    L.lsb200_sort_pairs_static.argtypes = [vp, vp, vp, vp, vp, C.c_int64]
    _libs[lab] = L
    return L


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def reachable_nodes(nodes, root):
    """Copy of a debug_copy_bvh node array with every record that no child reference leads to zeroed.  The fused tree
    build does not write the records of nodes whose whole range fits a leaf (their parents reference them as leaves), so
    those rows hold whatever the allocation held: tree comparisons look at the reachable records only."""
    child = nodes[:, 12:14].view(np.int32)
    seen = np.zeros(nodes.shape[0], bool)
    frontier = np.zeros(1, np.int64)
    while frontier.size:
        seen[frontier] = True
        c = child[frontier].ravel().astype(np.int64)
        frontier = c[c >= 0] - root
    out = nodes.copy()
    out[~seen] = 0
    return out


class Lsb200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"lsb200 error {code}: {msg}")
        self.code = code


class Context:
    """Thin object wrapper: one lsb200_ctx.  Methods return the ABI's status codes where the reference's
    interface does (add/remove/commit/trace) and raise Lsb200Error on CUDA failures."""

    def __init__(self, device=0, frame_mode=FRAME_SENSOR, stream=None, lab=False, _handle=None):
        self._L = lib(lab)
        self._lab = lab
        self._keep = {}
        if _handle is not None:              # adopt a context created by the library (lsb200_clone_scene)
            self._h = _handle
            return
        h = C.c_void_p()
        rc = self._L.lsb200_create(int(device), C.byref(h))
        if rc != 0:
            raise Lsb200Error(rc, "lsb200_create failed (no CUDA device? this library has no CPU path)")
        self._h = h
        if frame_mode != FRAME_SENSOR:
            self._chk(self._L.lsb200_set_frame_mode(h, frame_mode))
        if stream is not None:
            self.set_stream(stream)

    # -- helpers
    def _chk(self, rc):
        if rc < 0:
            raise Lsb200Error(rc, self.last_error())
        return rc

    def last_error(self):
        return self._L.lsb200_last_error(self._h).decode()

    def close(self):
        if self._h:
            self._L.lsb200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream_handle):
        self._chk(self._L.lsb200_set_stream(self._h, C.c_void_p(int(cuda_stream_handle) if cuda_stream_handle else None)))

    def set_option(self, opt, value):
        self._chk(self._L.lsb200_set_option(self._h, opt, int(value)))

    def set_frame_mode(self, mode):
        self._chk(self._L.lsb200_set_frame_mode(self._h, mode))

    def set_sensor(self, elev_deg, az_begin, az_end, H, q_wxyz=(1, 0, 0, 0), t_xyz=(0, 0, 0), index=0):
        e = np.ascontiguousarray(elev_deg, np.float32)
        q = np.ascontiguousarray(q_wxyz, np.float32)
        t = np.ascontiguousarray(t_xyz, np.float32)
        self._chk(self._L.lsb200_set_sensor(self._h, index, _fp(e), e.size, float(np.float32(az_begin)),
                                            float(np.float32(az_end)), int(H), _fp(q), _fp(t)))

    def set_sensor_limits(self, max_range=0.0, az_min_deg=-180.0, az_max_deg=180.0, index=0):
        return self._L.lsb200_set_sensor_limits(self._h, index, float(max_range), float(az_min_deg), float(az_max_deg))

    def set_sensor_pose(self, q_wxyz, t_xyz, index=0):
        q = np.ascontiguousarray(q_wxyz, np.float32)
        t = np.ascontiguousarray(t_xyz, np.float32)
        return self._L.lsb200_set_sensor_pose(self._h, index, _fp(q), _fp(t))

    def set_sensor_intensity_filter(self, min_intensity, index=0):
        return self._L.lsb200_set_sensor_intensity_filter(self._h, index, float(min_intensity))

    def set_log_callback(self, fn):
        """fn(level, message) or None.  The ctypes thunk is kept alive by the context."""
        if fn is None:
            self._keep["log"] = LOG_FN(0)
        else:
            self._keep["log"] = LOG_FN(lambda level, msg, user: fn(level, msg.decode(errors="replace")))
        return self._L.lsb200_set_log_callback(self._h, self._keep["log"], None)

    def clone(self, device):
        """A replica of this context's scene on another CUDA device (lsb200_clone_scene)."""
        h = C.c_void_p()
        rc = self._L.lsb200_clone_scene(self._h, int(device), C.byref(h))
        if rc != 0:
            raise Lsb200Error(rc, self.last_error())
        return Context(lab=self._lab, _handle=h)

    def alloc_host(self, nbytes):
        """Pinned host memory on the GPU's NUMA node as a uint8 numpy array (free with free_host)."""
        p = C.c_void_p()
        self._chk(self._L.lsb200_alloc_host(self._h, int(nbytes), C.byref(p)))
        arr = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(int(nbytes),))
        self._keep.setdefault("host", {})[arr.ctypes.data] = p
        return arr

    def free_host(self, arr):
        p = self._keep.get("host", {}).pop(arr.ctypes.data, None)
        if p is not None:
            self._L.lsb200_free_host(self._h, p)

    def numa_node(self):
        return self._L.lsb200_device_numa_node(self._h)

    def fetch_intensity(self, sensor=0):
        out = np.empty(self.ray_count(sensor), np.float32)
        rc = self._L.lsb200_fetch_intensity(self._h, sensor, _fp(out))
        return rc, out

    def expand_range_image(self, t, flags=0, sensor=0):
        """Host-side expansion of a range image into point records (ray-id order)."""
        t = np.ascontiguousarray(t, np.float32)
        rec = 16 if flags & TRACE_XYZI16 else 32
        out = np.empty((t.size, rec), np.uint8)
        n = C.c_uint32(0)
        rc = self._L.lsb200_expand_range_image(self._h, sensor, _fp(t), out.ctypes.data, out.shape[0], C.byref(n), flags)
        return rc, out[:n.value]

    def ray_count(self, index=0):
        return self._L.lsb200_get_ray_count(self._h, index)

    # -- ITracer-shaped calls (return codes like the reference)
    def add_geometry(self, name, geometry_type, n_vertices, n_elements):
        return self._L.lsb200_add_geometry(self._h, name.encode(), geometry_type, n_vertices, n_elements)

    def remove_geometry(self, name):
        return self._L.lsb200_remove_geometry(self._h, name.encode())

    def geometry_count(self):
        return self._L.lsb200_get_geometry_count(self._h)

    def geometry_id(self, name):
        return self._L.lsb200_get_geometry_id(self._h, name.encode())

    def update_geometry(self, name, verts=None, elems=None, affine=None, stride=None):
        """verts: (n,3) float32 array or a uint8 blob with `stride`; elems: (m,3|4) uint32; affine: 12 floats."""
        A = np.ascontiguousarray(affine if affine is not None else [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)
        vptr, eptr = None, None
        if verts is not None:
            if stride is None:
                verts = np.ascontiguousarray(verts, np.float32).reshape(-1, 3)
                stride = 12
            else:
                verts = np.ascontiguousarray(verts, np.uint8)
            vptr = verts.ctypes.data
        if elems is not None:
            elems = np.ascontiguousarray(elems, np.uint32)
            eptr = elems.ctypes.data
        self._keep[name] = (verts, elems)      # pinned callers may hand memory that is read asynchronously
        return self._L.lsb200_update_geometry(self._h, name.encode(), vptr, int(stride or 0), eptr, _fp(A))

    def update_geometry_euler(self, name, verts, elems, translation, euler_xyz, stride=None):
        A = affine_from_components(translation, euler_xyz)
        return self.update_geometry(name, verts, elems, A, stride)

    def update_transform(self, name, affine):
        A = np.ascontiguousarray(affine, np.float32)
        return self._L.lsb200_update_transform(self._h, name.encode(), _fp(A))

    def commit_scene(self, mode=COMMIT_AUTO):
        return self._L.lsb200_commit_scene(self._h, mode)

    def trace_scene(self, frame_index=0, flags=0, sensor=0, out=None):
        """Returns (status, points) with points a (n, 32|16) uint8 array (host)."""
        R = max(self.ray_count(sensor), 0)
        rec = 4 if flags & TRACE_RANGE_IMAGE else (16 if flags & TRACE_XYZI16 else 32)
        if out is None:
            out = np.empty((R, rec), np.uint8)
        n = C.c_uint32(0)
        rc = self._L.lsb200_trace_scene(self._h, sensor, frame_index, out.ctypes.data, out.shape[0], C.byref(n), flags)
        return rc, out[:n.value]

    def trace_device(self, sensor=0, flags=0):
        return self._L.lsb200_trace_device(self._h, sensor, flags)

    def trace_batch(self, first, count, flags=0):
        return self._L.lsb200_trace_batch(self._h, first, count, flags)

    def fetch_points(self, sensor=0, flags=0, out=None):
        R = max(self.ray_count(sensor), 0)
        rec = 4 if flags & TRACE_RANGE_IMAGE else (16 if flags & TRACE_XYZI16 else 32)
        if out is None:
            out = np.empty((R, rec), np.uint8)
        n = C.c_uint32(0)
        rc = self._L.lsb200_fetch_points(self._h, sensor, out.ctypes.data, out.shape[0], C.byref(n), flags)
        return rc, out[:n.value]

    def fetch_batch(self, first, count, out_ptrs, caps, flags=0):
        """out_ptrs: host addresses (ints), caps: capacities in records.  Returns (status, counts[count])."""
        outs = (C.c_void_p * count)(*[C.c_void_p(int(p)) for p in out_ptrs])
        cps = (C.c_size_t * count)(*[int(x) for x in caps])
        n = np.zeros(count, np.uint32)
        rc = self._L.lsb200_fetch_batch(self._h, first, count, outs, cps, n.ctypes.data_as(C.POINTER(C.c_uint32)), flags)
        return rc, n

    def run_frames(self, affines, out_ptrs, caps, first=0, count=1, commit_mode=COMMIT_AUTO, flags=0, verts=None):
        """Pipelined batch of frames.  affines: (F, G, 12) float32; out_ptrs: F lists of `count` host addresses;
        caps: `count` capacities (records); verts: optional list of G host addresses re-uploaded every frame.
        Returns (status, counts[F, count])."""
        A = np.ascontiguousarray(affines, np.float32)
        F = A.shape[0]
        counts = np.zeros((F, count), np.uint32)
        cps = (C.c_size_t * count)(*[int(x) for x in caps])
        vv = None
        if verts is not None:
            vv = (C.c_void_p * len(verts))(*[C.c_void_p(int(p)) if p else None for p in verts])
        frames = (Frame * F)()
        keep = []
        for f in range(F):
            outs = (C.c_void_p * count)(*[C.c_void_p(int(p)) for p in out_ptrs[f]])
            keep.append(outs)
            frames[f].affines = A[f].ctypes.data_as(C.POINTER(C.c_float))
            frames[f].verts = C.cast(vv, C.POINTER(C.c_void_p)) if vv is not None else None
            frames[f].out_points = C.cast(outs, C.POINTER(C.c_void_p))
            frames[f].capacities = C.cast(cps, C.POINTER(C.c_size_t))
            frames[f].n_pts = counts[f].ctypes.data_as(C.POINTER(C.c_uint32))
        rc = self._L.lsb200_run_frames(self._h, frames, F, first, count, commit_mode, flags)
        return rc, counts

    def trace_raw(self, sensor=0):
        R = self.ray_count(sensor)
        t, tri = np.empty(R, np.float32), np.empty(R, np.int32)
        rc = self._L.lsb200_trace_raw(self._h, sensor, _fp(t), tri.ctypes.data_as(C.POINTER(C.c_int32)))
        return rc, t, tri

    def fetch_second_returns(self, sensor=0):
        """(status, t2, tri2) of the last trace of `sensor` with OPT_RETURNS = 2."""
        R = self.ray_count(sensor)
        t, tri = np.empty(R, np.float32), np.empty(R, np.int32)
        rc = self._L.lsb200_fetch_second_returns(self._h, sensor, _fp(t), tri.ctypes.data_as(C.POINTER(C.c_int32)))
        return rc, t, tri

    def synchronize(self):
        self._chk(self._L.lsb200_synchronize(self._h))

    def init_message(self, frame_index):
        h = CloudHeader()
        self._chk(self._L.lsb200_init_message(self._h, frame_index, C.byref(h)))
        return h

    def stats(self):
        s = Stats()
        self._chk(self._L.lsb200_get_stats(self._h, C.byref(s)))
        return s

    def trace_counters(self):
        """Totals of the last OPT_TRACE_STATS launch as a dict."""
        out = (C.c_uint64 * 152)()
        self._chk(self._L.lsb200_get_trace_counters(self._h, out, 152))
        names = ("rays", "nodes", "leaves", "tris", "units_uniform", "units_mixed", "max_stack", "hits", "culled_pops", "chain_tests",
                 "tile_visits", "tile_entries", "tile_fallbacks", "tile_empty", "super_visits")
        d = {k: int(out[i]) for i, k in enumerate(names)}
        # per-unit cost histogram: bin b = units that took [2^b, 2^(b+1)) SM cycles
        for j, key in enumerate(("cyc_units", "cyc_sum", "cyc_nodes", "cyc_list")):
            d[key] = [int(out[15 + 24 * j + b]) for b in range(24)]
        d["heavy_cycles_by_class"] = [int(out[111 + o]) for o in range(9)]
        d["heavy_cycles_by_band"] = [int(out[120 + b]) for b in range(32)]
        return d

    def reset_counters(self):
        self._L.lsb200_reset_counters(self._h)

    def debug_copy_bvh(self):
        s = self.stats()
        nodes = np.empty((max(int(s.n_nodes), 1), 16), np.float32)
        ids = np.empty(int(s.n_triangles), np.int32)
        root = C.c_int32(0)
        self._chk(self._L.lsb200_debug_copy_bvh(self._h, _fp(nodes), nodes.shape[0], ids.ctypes.data_as(C.POINTER(C.c_int32)),
                                                ids.size, C.byref(root)))
        return nodes, ids, root.value

    def debug_sort_pairs(self, keys, vals):
        k = np.ascontiguousarray(keys, np.uint32).copy()
        v = np.ascontiguousarray(vals, np.uint32).copy()
        self._chk(self._L.lsb200_debug_sort_pairs(self._h, k.ctypes.data_as(C.POINTER(C.c_uint32)),
                                                  v.ctypes.data_as(C.POINTER(C.c_uint32)), k.size))
        return k, v

    def sort_pairs_static(self, d_keys, d_vals, d_keys_tmp, d_vals_tmp, n):
        """lsb200_sort_pairs_static: DEVICE addresses (ints) of the caller's arrays; enqueued on the context's stream,
        not synchronised; allocates nothing.  Returns the ABI's status code (ERR_CAPACITY beyond SORT_STATIC_MAX_PAIRS)."""
        return self._L.lsb200_sort_pairs_static(self._h, C.c_void_p(d_keys), C.c_void_p(d_vals), C.c_void_p(d_keys_tmp),
                                                C.c_void_p(d_vals_tmp), int(n))

    def debug_copy_vertices(self):
        s = self.stats()
        out = np.empty((int(s.n_vertices), 3), np.float32)
        self._chk(self._L.lsb200_debug_copy_vertices(self._h, _fp(out), out.shape[0]))
        return out


def _build_frames(affines, out_ptrs, caps, count, verts):
    A = np.ascontiguousarray(affines, np.float32)
    F = A.shape[0]
    counts = np.zeros((F, count), np.uint32)
    cps = (C.c_size_t * count)(*[int(x) for x in caps])
    vv = None
    if verts is not None:
        vv = (C.c_void_p * len(verts))(*[C.c_void_p(int(p)) if p else None for p in verts])
    frames = (Frame * F)()
    keep = [A, cps, vv]
    for f in range(F):
        outs = (C.c_void_p * count)(*[C.c_void_p(int(p)) for p in out_ptrs[f]])
        keep.append(outs)
        frames[f].affines = A[f].ctypes.data_as(C.POINTER(C.c_float))
        frames[f].verts = C.cast(vv, C.POINTER(C.c_void_p)) if vv is not None else None
        frames[f].out_points = C.cast(outs, C.POINTER(C.c_void_p))
        frames[f].capacities = C.cast(cps, C.POINTER(C.c_size_t))
        frames[f].n_pts = counts[f].ctypes.data_as(C.POINTER(C.c_uint32))
    return frames, counts, keep


def run_batch(contexts, affines, out_ptrs, caps, first=0, count=1, commit_mode=COMMIT_AUTO, flags=0, sharding=SHARD_FRAMES):
    """lsb200_run_batch: the frames on several contexts (one per GPU) at once, one host thread each inside the library.
    Returns (status, counts[F, count])."""
    frames, counts, keep = _build_frames(affines, out_ptrs, caps, count, None)
    hs = (C.c_void_p * len(contexts))(*[c._h for c in contexts])
    rc = contexts[0]._L.lsb200_run_batch(hs, len(contexts), frames, len(frames), first, count, commit_mode, flags, sharding)
    return rc, counts


def affine_from_components(translation, euler_xyz):
    t = np.ascontiguousarray(translation, np.float32)
    r = np.ascontiguousarray(euler_xyz, np.float32)
    A = np.empty(12, np.float32)
    lib().lsb200_affine_from_components(_fp(t), _fp(r), _fp(A))
    return A
