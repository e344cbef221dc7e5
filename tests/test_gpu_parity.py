"""GPU parity: the CUDA path, called through the C ABI (include/lsb200.h), against the CPU oracle.

Bar (BASELINE.json north_star): hit/miss mask and triangle id bit-exact, range within 1e-4 relative.  The CUDA
kernels use the oracle's operation order without FMA contraction, so these tests demand MORE: t bit-exact too.
Reference pins: 1668 / 1781 / 0 points (SRC/test/EmbreeTracer_test.cpp:134, SRC/test/OptixTracer_test.cpp:168,307).
"""
import numpy as np
import pytest

from conftest import xt32_sensor_args
from lidarshooter_b200 import capi, synth
from oracle import lso

pytestmark = pytest.mark.gpu

IDENT = lso.IDENTITY_AFFINE


def make_ctx(uid="0000", frame_mode=capi.FRAME_SENSOR, opts=None, lab=False):
    ctx = capi.Context(0, frame_mode, lab=lab)
    for k, v in (opts or {}).items():
        ctx.set_option(k, v)
    elev, b, e, H, q, t = xt32_sensor_args(uid)
    ctx.set_sensor(elev, b, e, H, q, t)
    return ctx


def add_mesh(ctx, name, v, t, A=IDENT):
    gid = ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0])
    assert gid >= 0
    assert ctx.update_geometry(name, v, t, A) == 0
    return gid


def sort_rows(p):
    """Canonical order of a point set: lexicographic over the raw 32-bit words."""
    w = np.ascontiguousarray(p).view(np.uint32).reshape(p.shape[0], -1)
    return p[np.lexsort(w.T[::-1])]


@pytest.mark.parametrize("small_trace_units", [512, 0])     # one thread per ray (k_trace_small, the default at this size) / the persistent queue kernel
@pytest.mark.parametrize("uid,n_ground,n_both", [("0000", 1668, 1781), ("0001", 1633, 1769)])
def test_xt32_reference_counts_and_golden(meshes, golden, uid, n_ground, n_both, small_trace_units):
    ctx = make_ctx(uid)
    ctx.set_option(capi.OPT_SMALL_TRACE_UNITS, small_trace_units)
    add_mesh(ctx, "ground", meshes["ground_v"], meshes["ground_t"])
    assert ctx.commit_scene() == 0
    rc, pts = ctx.trace_scene(0, capi.TRACE_ORDERED)
    assert rc == 0 and pts.shape == (n_ground, 32)
    rc, t, tri = ctx.trace_raw()
    assert rc == 0
    assert np.array_equal(tri, golden[f"l{uid}_ground_tri"])
    assert np.array_equal(t, golden[f"l{uid}_ground_t"])

    add_mesh(ctx, "ben", meshes["ben_v"], meshes["ben_t"])
    assert ctx.commit_scene() == 0
    rc, t, tri = ctx.trace_raw()
    assert np.array_equal(tri, golden[f"l{uid}_ground_ben_tri"])
    assert np.array_equal(t, golden[f"l{uid}_ground_ben_t"])
    rc, pts = ctx.trace_scene(1, capi.TRACE_ORDERED)
    assert rc == 0 and pts.shape[0] == n_both

    # oracle on the same inputs: packed points byte-identical in ray order; default (atomic) order is a permutation
    s = lso.Sensor(*xt32_sensor_args(uid))
    ref = lso.scan(s, [(meshes["ground_v"], meshes["ground_t"], IDENT), (meshes["ben_v"], meshes["ben_t"], IDENT)])
    assert np.array_equal(pts, ref["points"])
    rc, pts_u = ctx.trace_scene(2, 0)
    assert rc == 0 and pts_u.shape[0] == n_both
    assert np.array_equal(sort_rows(pts_u), sort_rows(ref["points"]))
    # Embree-path quirk: ring 0 (SRC/EmbreeTracer.cpp:333-334)
    rc, pts0 = ctx.trace_scene(3, capi.TRACE_ORDERED | capi.TRACE_RING_ZERO)
    ref0 = lso.scan(s, [(meshes["ground_v"], meshes["ground_t"], IDENT), (meshes["ben_v"], meshes["ben_t"], IDENT)], ring_mode=1)
    assert np.array_equal(pts0, ref0["points"])
    # 16-byte XYZI records
    rc, p16 = ctx.trace_scene(4, capi.TRACE_ORDERED | capi.TRACE_XYZI16)
    f = ref["points"].view(np.float32).reshape(-1, 8)
    assert np.array_equal(p16.view(np.float32).reshape(-1, 4), f[:, [0, 1, 2, 4]])

    # remove ben -> back to the ground-only count (SRC/test/OptixTracer_test.cpp:217-290)
    assert ctx.remove_geometry("ben") == 1
    assert ctx.geometry_count() == 1
    rc, pts = ctx.trace_scene(5, capi.TRACE_ORDERED)
    assert rc == 0 and pts.shape[0] == n_ground
    # remove the last one -> empty scene: -1 and an empty cloud (SRC/test/OptixTracer_test.cpp:297-310)
    assert ctx.remove_geometry("ground") == 0
    assert ctx.geometry_count() == 0
    assert ctx.commit_scene() == -1
    rc, pts = ctx.trace_scene(6, 0)
    assert rc == -1 and pts.shape[0] == 0
    assert ctx.remove_geometry("nope") == -1
    ctx.close()


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4, 5, 6, 7])
@pytest.mark.parametrize("leaf,tile", [(1, 8), (4, 32), (8, 2), (2, 16)])
def test_traversal_variants_agree(meshes, golden, mode, leaf, tile):
    ctx = make_ctx("0000", opts={capi.OPT_TRAVERSAL: mode, capi.OPT_LEAF_SIZE: leaf, capi.OPT_TOP_NODES: 256,
                                 capi.OPT_TILE_WIDTH: tile}, lab=True)
    add_mesh(ctx, "ground", meshes["ground_v"], meshes["ground_t"])
    add_mesh(ctx, "ben", meshes["ben_v"], meshes["ben_t"])
    assert ctx.commit_scene() == 0
    rc, t, tri = ctx.trace_raw()
    assert rc == 0
    assert np.array_equal(tri, golden["l0000_ground_ben_tri"])
    assert np.array_equal(t, golden["l0000_ground_ben_t"])
    ctx.close()


def test_transform_bit_exact_and_refit(meshes):
    """updateGeometry with translation + Euler angles (SRC/MeshTransformer.cpp:467-477), rebuild vs refit."""
    s = lso.Sensor(*xt32_sensor_args("0000"))
    ctx = make_ctx("0000")
    add_mesh(ctx, "ground", meshes["ground_v"], meshes["ground_t"])
    add_mesh(ctx, "ben", meshes["ben_v"], meshes["ben_t"])
    poses = synth.trajectory_poses([(1.0, (0, 5, 0), (0, 0, 1)), (0.5, (1, 0, 0), (0.3, -0.2, -1))])
    for k, (lin, ang) in enumerate(poses):
        A = lso.affine_from_components(lin, ang)
        assert np.array_equal(A, capi.affine_from_components(lin, ang))
        assert ctx.update_geometry_euler("ben", None, None, lin, ang) == 0
        assert ctx.commit_scene(capi.COMMIT_AUTO) == 0
        st = ctx.stats()
        assert st.last_commit_was_rebuild == (1 if k == 0 else 0)
        ref = lso.scan(s, [(meshes["ground_v"], meshes["ground_t"], IDENT), (meshes["ben_v"], meshes["ben_t"], A)])
        if k in (0, 7, 14):
            assert np.array_equal(ctx.debug_copy_vertices(), ref["verts"])
        rc, t, tri = ctx.trace_raw()
        assert rc == 0
        assert np.array_equal(tri, ref["tri"]) and np.array_equal(t, ref["t"])
    # a forced rebuild gives the same answers as the refitted tree
    assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0
    rc, t2, tri2 = ctx.trace_raw()
    assert np.array_equal(tri2, tri) and np.array_equal(t2, t)
    ctx.close()


def test_strided_vertex_blob_and_quads():
    """pcl::PCLPointCloud2 blobs carry point_step bytes per vertex; quads are split (v0 v1 v3), (v2 v3 v1)."""
    rng = np.random.default_rng(3)
    n = 40
    xs, ys = np.meshgrid(np.linspace(-30, 30, n + 1), np.linspace(-30, 30, n + 1), indexing="xy")
    v = np.stack([xs, ys, rng.normal(-2.0, 0.3, xs.shape)], -1).reshape(-1, 3).astype(np.float32)
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="xy")
    v00 = (j * (n + 1) + i).astype(np.uint32)
    quads = np.stack([v00, v00 + 1, v00 + n + 2, v00 + n + 1], -1).reshape(-1, 4)
    tris = np.stack([quads[:, [0, 1, 3]], quads[:, [2, 3, 1]]], 1).reshape(-1, 3)
    blob = np.zeros((v.shape[0], 8), np.float32)
    blob[:, :3] = v
    blob[:, 3:] = 7.0
    elev, b, e, H = synth.XT32_ELEV, 0.0, 360.0, 150
    ctx = capi.Context(0)
    ctx.set_sensor(elev, b, e, H, (1, 0, 0, 0), (0, 0, 0))
    assert ctx.add_geometry("q", capi.GEOMETRY_QUAD, v.shape[0], quads.shape[0]) == 0
    assert ctx.add_geometry("bad", 24, 3, 1) < 0                      # unsupported RTCGeometryType
    assert ctx.update_geometry("q", blob.view(np.uint8), quads, IDENT, stride=32) == 0
    assert ctx.commit_scene() == 0
    rc, t, tri = ctx.trace_raw()
    s = lso.Sensor(elev, b, e, H)
    ref = lso.scan(s, [(v, tris, IDENT)])
    assert (ref["tri"] >= 0).sum() > 500
    assert np.array_equal(tri, ref["tri"]) and np.array_equal(t, ref["t"])
    ctx.close()


def test_world_frame_multi_sensor(meshes):
    """World-frame mode: one BVH, rays moved (o = T, d = R d_s); bit-exact against the oracle in the same mode."""
    ctx = capi.Context(0, capi.FRAME_WORLD)
    sens = []
    for k, uid in enumerate(("0000", "0001")):
        args = xt32_sensor_args(uid)
        ctx.set_sensor(*args, index=k)
        sens.append(lso.Sensor(*args))
    add_mesh(ctx, "ground", meshes["ground_v"], meshes["ground_t"])
    add_mesh(ctx, "ben", meshes["ben_v"], meshes["ben_t"])
    assert ctx.commit_scene() == 0
    assert ctx.trace_batch(0, 2, capi.TRACE_ORDERED) == 0
    for k in range(2):
        ref = lso.scan(sens[k], [(meshes["ground_v"], meshes["ground_t"], IDENT), (meshes["ben_v"], meshes["ben_t"], IDENT)], mode=1)
        rc, pts = ctx.fetch_points(k, capi.TRACE_ORDERED)
        assert rc == 0
        assert np.array_equal(pts, ref["points"])
        rc, t, tri = ctx.trace_raw(k)
        assert np.array_equal(tri, ref["tri"]) and np.array_equal(t, ref["t"])
    ctx.close()


@pytest.mark.parametrize("flags,vert_upload", [(capi.TRACE_ORDERED, False), (0, True), (capi.TRACE_ORDERED | capi.TRACE_XYZI16, True)])
def test_run_frames_pipelined_matches_per_frame_calls(meshes, flags, vert_upload):
    """lsb200_run_frames (two output slots, D2H of frame f overlapped with frame f+1) == the per-frame
    updateGeometry / commitScene / traceScene sequence, and == the oracle, for two sensors along a trajectory."""
    ctx = capi.Context(0, capi.FRAME_WORLD)
    sens = []
    for k, uid in enumerate(("0000", "0001")):
        args = xt32_sensor_args(uid)
        ctx.set_sensor(*args, index=k)
        sens.append(lso.Sensor(*args))
    gv, gt, bv, bt = meshes["ground_v"], meshes["ground_t"], meshes["ben_v"], meshes["ben_t"]
    add_mesh(ctx, "ground", gv, gt)
    add_mesh(ctx, "ben", bv, bt)
    poses = synth.trajectory_poses([(0.7, (0, 5, 0), (0, 0, 1))])
    F = len(poses)
    A = np.zeros((F, 2, 12), np.float32)
    for f, (lin, ang) in enumerate(poses):
        A[f, 0] = IDENT
        A[f, 1] = lso.affine_from_components(lin, ang)
    rec = 16 if flags & capi.TRACE_XYZI16 else 32
    R = ctx.ray_count(0)
    outs = [[np.zeros((R, rec), np.uint8) for _ in range(2)] for _ in range(F)]
    ptrs = [[o.ctypes.data for o in row] for row in outs]
    gvc, bvc = np.ascontiguousarray(gv, np.float32), np.ascontiguousarray(bv, np.float32)
    verts = [gvc.ctypes.data, bvc.ctypes.data] if vert_upload else None
    rc, counts = ctx.run_frames(A, ptrs, [R, R], first=0, count=2, commit_mode=capi.COMMIT_AUTO, flags=flags, verts=verts)
    assert rc == 0, ctx.last_error()
    for f in range(F):
        for k in range(2):
            ref = lso.scan(sens[k], [(gv, gt, IDENT), (bv, bt, A[f, 1])], mode=1)
            want = ref["points"]
            if rec == 16:
                want = np.ascontiguousarray(want.view(np.float32).reshape(-1, 8)[:, [0, 1, 2, 4]]).view(np.uint8).reshape(-1, 16)
            got = outs[f][k][:counts[f, k]]
            assert counts[f, k] == want.shape[0]
            if flags & capi.TRACE_ORDERED:
                assert np.array_equal(got, want)
            else:
                assert np.array_equal(sort_rows(got), sort_rows(want))
    # capacity error is reported, counts still filled
    rc, counts2 = ctx.run_frames(A[:2], ptrs[:2], [10, 10], first=0, count=2, flags=flags)
    assert rc == capi.ERR_CAPACITY and np.array_equal(counts2, counts[:2])
    ctx.close()


@pytest.mark.parametrize("n_tris", [1, 2, 3, 100, 5651, 8192, 16384, 16385])
def test_small_build_equals_pipeline(n_tris):
    """The fused single-CTA commit of small scenes (k_build_small) builds the SAME tree as the multi-kernel pipeline:
    node array, leaf order and per-ray results are bit-identical; refit after a transform change too."""
    rng = np.random.default_rng(7 * n_tris + 1)
    c = rng.uniform(-20, 20, (n_tris, 1, 3)).astype(np.float32)
    c[: n_tris // 3] = c[0]                                  # many equal Morton codes: ties resolved by index
    v = (c + rng.normal(0, 0.5, (n_tris, 3, 3)).astype(np.float32)).reshape(-1, 3)
    tr = np.arange(3 * n_tris, dtype=np.uint32).reshape(-1, 3)
    elev = np.linspace(60, -60, 24).astype(np.float32)
    A2 = lso.affine_from_components([0.5, -0.25, 0.125], [0.1, -0.2, 0.3])
    out = []
    for small in (1, 0):
        ctx = capi.Context(0)
        ctx.set_option(capi.OPT_SMALL_BUILD, small)
        ctx.set_sensor(elev, 0.0, 360.0, 90, (0.9, 0.1, -0.2, 0.3), (1.0, -2.0, 0.5))
        add_mesh(ctx, "soup", v, tr)
        assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0, ctx.last_error()
        nodes, ids, root = ctx.debug_copy_bvh()
        rc, t, tri = ctx.trace_raw()
        assert ctx.update_transform("soup", A2) == 0
        assert ctx.commit_scene(capi.COMMIT_REFIT) == 0 and ctx.stats().last_commit_was_rebuild == 0
        nodes2, ids2, root2 = ctx.debug_copy_bvh()
        rc, t2, tri2 = ctx.trace_raw()
        out.append((capi.reachable_nodes(nodes, root), ids.copy(), root, t, tri, capi.reachable_nodes(nodes2, root2), ids2.copy(), t2, tri2))
        ctx.close()
    a, b = out
    names = ("nodes", "leaf ids", "root", "t", "tri", "nodes after refit", "leaf ids after refit", "t after refit", "tri after refit")
    for name, x, y in zip(names, a, b):
        assert np.array_equal(np.atleast_1d(x).view(np.uint8), np.atleast_1d(y).view(np.uint8)), name
    assert n_tris < 100 or (a[4] >= 0).sum() > 0


@pytest.mark.parametrize("lists", ["default", "tile", "tile+super"])
@pytest.mark.parametrize("n_tris", [1, 2, 3, 5, 1000, 20000])
def test_random_soup_vs_brute(n_tris, lists):
    rng = np.random.default_rng(n_tris)
    c = rng.uniform(-20, 20, (n_tris, 1, 3)).astype(np.float32)
    v = (c + rng.normal(0, 0.7, (n_tris, 3, 3)).astype(np.float32)).reshape(-1, 3)
    tr = np.arange(3 * n_tris, dtype=np.uint32).reshape(-1, 3)
    elev = np.linspace(80, -80, 64).astype(np.float32)
    ctx = capi.Context(0, lab=True)
    if lists != "default":          # per-tile entry lists forced on for this small table (sensor inside the soup, tiny trees)
        ctx.set_option(capi.OPT_TILE_MIN_UNITS, 0)
        ctx.set_option(capi.OPT_SUPER_TILES, 1 if lists == "tile+super" else 0)
    ctx.set_sensor(elev, 0.0, 360.0, 181, (1, 0, 0, 0), (0, 0, 0))
    add_mesh(ctx, "soup", v, tr)
    assert ctx.commit_scene() == 0
    rc, t, tri = ctx.trace_raw()
    s = lso.Sensor(elev, 0.0, 360.0, 181)
    ref = lso.scan(s, [(v, tr, IDENT)], use_bvh=False)
    assert np.array_equal(tri, ref["tri"]) and np.array_equal(t, ref["t"])
    ctx.close()


def _axis_aligned_room(half, z0, z1):
    """Closed axis-aligned box room around the origin (12 triangles) + a flat plate in the z = 0 plane (2 triangles):
    every node box has at least one zero-thickness axis somewhere down the tree."""
    x0, x1, y0, y1 = -half, half, -half, half
    c = np.array([[x0, y0, z0], [x1, y0, z0], [x1, y1, z0], [x0, y1, z0], [x0, y0, z1], [x1, y0, z1], [x1, y1, z1], [x0, y1, z1],
                  [1, -1, 0], [3, -1, 0], [3, 1, 0], [1, 1, 0]], np.float32)
    q = [(0, 1, 2, 3), (4, 5, 6, 7), (0, 1, 5, 4), (1, 2, 6, 5), (2, 3, 7, 6), (3, 0, 4, 7), (8, 9, 10, 11)]
    tr = np.array([t for a, b, c_, d in q for t in ((a, b, c_), (a, c_, d))], np.uint32)
    return c, tr


@pytest.mark.parametrize("lists", ["default", "tile", "tile+super"])
@pytest.mark.parametrize("origin", [(0.0, 0.0, 0.0), (0.0, 0.0, 0.5), (2.0, 0.0, 0.0), (1e4, -2e4, 50.0)])
def test_axis_parallel_rays_and_flat_boxes(origin, lists):
    """Directions with exactly-zero (or rounding-noise) components -- elevation 0 / +-90 deg, azimuth steps of 45 deg --
    against axis-aligned walls and a plate in the sensor's own horizontal plane (rays IN the plane of a triangle, rays
    along box faces, an origin on a triangle): 0 x inf in a slab test must not lose or invent a hit.  Brute-force
    oracle, bit-exact."""
    v, tr = _axis_aligned_room(8.0, -2.0, 4.0)
    v = (v + np.array(origin, np.float32)).astype(np.float32) if abs(origin[0]) > 100 else v
    pos = origin
    elev = np.array([90, 45, 1e-3, 0, -1e-3, -45, -90, 30], np.float32)
    ctx = capi.Context(0, capi.FRAME_WORLD, lab=True)
    if lists != "default":
        ctx.set_option(capi.OPT_TILE_MIN_UNITS, 0)
        ctx.set_option(capi.OPT_SUPER_TILES, 1 if lists == "tile+super" else 0)
    ctx.set_sensor(elev, 0.0, 360.0, 17, (1, 0, 0, 0), pos)
    add_mesh(ctx, "room", v, tr)
    assert ctx.commit_scene() == 0, ctx.last_error()
    rc, t, tri = ctx.trace_raw()
    assert rc == 0
    s = lso.Sensor(elev, 0.0, 360.0, 17, (1, 0, 0, 0), pos)
    ref = lso.scan(s, [(v, tr, IDENT)], mode=1, use_bvh=False)
    assert np.array_equal(tri, ref["tri"]) and np.array_equal(t, ref["t"])
    assert (tri >= 0).sum() > len(tri) // 2
    ctx.close()


def test_duplicate_and_degenerate_triangles():
    """Coincident triangles (equal t): the smaller global id wins; zero-area triangles never hit."""
    v = np.array([[5, -1, -1], [5, 1, -1], [5, 0, 1], [6, 0, 0], [6, 0, 0], [6, 0, 0]], np.float32)
    tr = np.array([[0, 1, 2], [3, 4, 5], [0, 1, 2], [2, 1, 0]], np.uint32)
    ctx = capi.Context(0, lab=True)
    ctx.set_sensor(np.array([0.0], np.float32), -10.0, 10.0, 21, (1, 0, 0, 0), (0, 0, 0))
    add_mesh(ctx, "m", v, tr)
    assert ctx.commit_scene() == 0
    rc, t, tri = ctx.trace_raw()
    ctx.set_option(capi.OPT_TILE_MIN_UNITS, 0)          # same answers through the per-tile entry lists
    ctx.set_option(capi.OPT_SUPER_TILES, 1)
    rc2, t2, tri2 = ctx.trace_raw()
    assert rc2 == 0 and np.array_equal(t2, t) and np.array_equal(tri2, tri)
    s = lso.Sensor(np.array([0.0], np.float32), -10.0, 10.0, 21)
    ref = lso.scan(s, [(v, tr, IDENT)], use_bvh=False)
    assert np.array_equal(tri, ref["tri"]) and np.array_equal(t, ref["t"])
    assert set(np.unique(tri)) <= {-1, 0}
    ctx.close()


@pytest.mark.parametrize("n", [0, 1, 255, 4096, 4097, 100_000, 3_000_001])
def test_radix_sort_pairs(n):
    rng = np.random.default_rng(n + 1)
    keys = rng.integers(0, 1 << 30, n, dtype=np.uint32)
    if n > 1000:
        keys[: n // 3] &= np.uint32(0x3FF)          # long runs of equal high digits + duplicates
    vals = np.arange(n, dtype=np.uint32)
    ctx = capi.Context(0)
    k, v = ctx.debug_sort_pairs(keys, vals)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(k, keys[order])
    assert np.array_equal(v, vals[order])           # stable
    ctx.close()


def test_full_32bit_keys_sort():
    rng = np.random.default_rng(9)
    keys = rng.integers(0, 1 << 32, 50_000, dtype=np.uint64).astype(np.uint32)
    keys[:100] = 0xFFFFFFFF
    vals = np.arange(keys.size, dtype=np.uint32)
    ctx = capi.Context(0)
    k, v = ctx.debug_sort_pairs(keys, vals)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(k, keys[order]) and np.array_equal(v, vals[order])
    ctx.close()


def test_bvh_structure(meshes):
    """Every triangle is referenced by exactly one leaf; child boxes contain their triangles."""
    ctx = make_ctx("0000", opts={capi.OPT_LEAF_SIZE: 4})
    add_mesh(ctx, "ground", meshes["ground_v"], meshes["ground_t"])
    add_mesh(ctx, "ben", meshes["ben_v"], meshes["ben_t"])
    assert ctx.commit_scene() == 0
    nodes, ids, root = ctx.debug_copy_bvh()
    nT = ids.size
    assert sorted(ids.tolist()) == list(range(nT))
    xv = ctx.debug_copy_vertices()
    s = lso.Sensor(*xt32_sensor_args("0000"))
    gv, gt = lso.concat_scene(s, [(meshes["ground_v"], meshes["ground_t"], IDENT), (meshes["ben_v"], meshes["ben_t"], IDENT)])
    assert np.array_equal(xv, gv)
    tri_lo = gv[gt].min(1)[ids]          # leaf order
    tri_hi = gv[gt].max(1)[ids]
    child = nodes[:, 12:14].view(np.int32)
    seen = np.zeros(nT, np.int32)
    stack = [0]
    visited = 0
    while stack:
        k = stack.pop()
        visited += 1
        n = nodes[k]
        boxes = [((n[0], n[2], n[8]), (n[1], n[3], n[9])), ((n[4], n[6], n[10]), (n[5], n[7], n[11]))]
        for sl in range(2):
            c = int(child[k, sl])
            lo, hi = np.array(boxes[sl][0]), np.array(boxes[sl][1])
            if c < 0:
                enc = ~c
                first, cnt = enc & ((1 << 28) - 1), (enc >> 28) + 1
                assert cnt <= 4
                seen[first:first + cnt] += 1
                assert (tri_lo[first:first + cnt] >= lo).all() and (tri_hi[first:first + cnt] <= hi).all()
            else:
                stack.append(c - root)
    assert (seen == 1).all()
    assert visited < nT
    ctx.close()


def test_heightfield_1m_tris_vs_oracle_bvh():
    """BASELINE config 3: 128 channels x 3601 azimuth steps against the 1M-triangle heightfield."""
    v, tr = synth.heightfield()
    assert tr.shape[0] == 1_000_000 and v.shape[0] == 501_501
    elev, b, e, H = synth.lidar_table()
    T = (0.0, 0.0, 4.0)
    ctx = capi.Context(0)
    ctx.set_sensor(elev, b, e, H, (1, 0, 0, 0), T)
    add_mesh(ctx, "terrain", v, tr)
    assert ctx.commit_scene() == 0
    rc, t, tri = ctx.trace_raw()
    assert rc == 0
    s = lso.Sensor(elev, b, e, H, (1, 0, 0, 0), T)
    ref = lso.scan(s, [(v, tr, IDENT)], use_bvh=True)
    n_hit = int((ref["tri"] >= 0).sum())
    assert n_hit > 200_000
    mism = int((tri != ref["tri"]).sum())
    assert mism == 0, f"{mism} rays differ in triangle id"
    assert np.array_equal(t, ref["t"])
    rc, pts = ctx.trace_scene(0, 0)
    assert pts.shape[0] == n_hit
    ctx.close()


@pytest.mark.parametrize("mode", [0, 1, 4, 5, 6, 7])
def test_small_city_world_frame_vs_oracle_bvh(mode):
    """BASELINE config 4 in miniature (same generator, 4 yawed sensors, world frame): every traversal variant
    bit-exact against the oracle's CPU BVH; rebuild == refit after the buildings move."""
    meshes = synth.city(n_boxes=380, ground_cells=200, extent=20 * synth.CITY_PITCH, n_groups=4, lots=20)
    poses = synth.sensor_lattice(2, 10 * synth.CITY_PITCH)
    elev, b, e, H = synth.lidar_table(32, 721)
    ctx = capi.Context(0, capi.FRAME_WORLD, lab=True)
    ctx.set_option(capi.OPT_TRAVERSAL, mode)
    for k, (q, t) in enumerate(poses):
        ctx.set_sensor(elev, b, e, H, q, t, index=k)
    A = [IDENT.copy() for _ in meshes]
    for gi, (name, v, tr) in enumerate(meshes):
        add_mesh(ctx, name, v, tr, A[gi])
    for it in range(2):
        assert ctx.commit_scene(capi.COMMIT_AUTO) == 0
        assert ctx.stats().last_commit_was_rebuild == (1 if it == 0 else 0)
        assert ctx.trace_batch(0, len(poses), capi.TRACE_ORDERED) == 0
        s0 = lso.Sensor(elev, b, e, H, poses[0][0], poses[0][1])
        verts, tris = lso.concat_scene(s0, [(v, tr, A[gi]) for gi, (_, v, tr) in enumerate(meshes)], mode=1)
        bvh = lso.Bvh(verts, tris)
        for k, (q, t) in enumerate(poses):
            s = lso.Sensor(elev, b, e, H, q, t)
            dirs_s = lso.ray_dirs(s.tables)
            t_ref, tri_ref = bvh.trace(s.T, lso.rotate_dirs(s.R, dirs_s))
            rc, tt, tri = ctx.trace_raw(k)
            assert rc == 0
            assert np.array_equal(tri, tri_ref) and np.array_equal(tt, t_ref)
            rc, pts = ctx.fetch_points(k, capi.TRACE_ORDERED)
            # trace_raw re-traced without points: fetch the ordered clouds of the batch again
        assert ctx.trace_batch(0, len(poses), capi.TRACE_ORDERED) == 0
        for k, (q, t) in enumerate(poses):
            s = lso.Sensor(elev, b, e, H, q, t)
            dirs_s = lso.ray_dirs(s.tables)
            t_ref, tri_ref = bvh.trace(s.T, lso.rotate_dirs(s.R, dirs_s))
            rc, pts = ctx.fetch_points(k, capi.TRACE_ORDERED)
            assert rc == 0 and np.array_equal(pts, lso.pack_points(dirs_s, t_ref, tri_ref, H))
        bvh.close()
        # move every building group a little and refit
        for gi in range(1, len(meshes)):
            A[gi] = lso.affine_from_components((0.3 * gi, -0.2 * gi, 0.0), (0.0, 0.0, 0.01 * gi))
            assert ctx.update_geometry(meshes[gi][0], None, None, A[gi]) == 0
    ctx.close()


@pytest.mark.parametrize("entries", [2, 3, 8, 12, 16])
@pytest.mark.parametrize("leaf,tile", [(1, 8), (2, 8), (4, 32), (8, 2), (2, 16)])
def test_tile_entry_lists_xt32_golden(meshes, golden, entries, leaf, tile):
    """Per-tile entry lists (k_tile_entries) forced on for the 150-tile XT-32 scan, sensor frame: every list size,
    leaf size and tile shape returns the golden per-ray (t, triangle id) bits."""
    ctx = make_ctx("0000", opts={capi.OPT_TRAVERSAL: 6, capi.OPT_LEAF_SIZE: leaf, capi.OPT_TILE_WIDTH: tile,
                                 capi.OPT_TILE_ENTRIES: entries, capi.OPT_TILE_MIN_UNITS: 0}, lab=True)
    add_mesh(ctx, "ground", meshes["ground_v"], meshes["ground_t"])
    add_mesh(ctx, "ben", meshes["ben_v"], meshes["ben_t"])
    assert ctx.commit_scene() == 0
    ctx.set_option(capi.OPT_TRACE_STATS, 1)
    rc, t, tri = ctx.trace_raw()
    assert rc == 0
    cnt = ctx.trace_counters()
    assert entries < 8 or tile > 8 or cnt["tile_entries"] > 0          # ... and its lists were used (wide tiles of an XT-32 are sign-mixed)
    assert np.array_equal(tri, golden["l0000_ground_ben_tri"])
    assert np.array_equal(t, golden["l0000_ground_ben_t"])
    ctx.set_option(capi.OPT_TRACE_STATS, 0)
    rc, pts = ctx.trace_scene(0, capi.TRACE_ORDERED)
    assert rc == 0 and pts.shape[0] == 1781
    ctx.close()


@pytest.mark.parametrize("super_tiles", [1, 0])
@pytest.mark.parametrize("entries,visits", [(2, 0), (6, 3), (10, 0), (16, 1000)])
def test_tile_entry_lists_small_city_vs_oracle_bvh(entries, visits, super_tiles):
    """World frame, 4 yawed sensors in a small city: tile lists of every size bit-exact against the oracle's CPU BVH,
    after a rebuild and after a refit (the lists are rebuilt with the tree), with a detection radius and an azimuth
    window (entries beyond the radius are dropped by the tile walk)."""
    meshes = synth.city(n_boxes=380, ground_cells=200, extent=20 * synth.CITY_PITCH, n_groups=4, lots=20)
    poses = synth.sensor_lattice(2, 10 * synth.CITY_PITCH)
    elev, b, e, H = synth.lidar_table(32, 721)
    ctx = capi.Context(0, capi.FRAME_WORLD, lab=True)
    ctx.set_option(capi.OPT_TILE_ENTRIES, entries)
    ctx.set_option(capi.OPT_TILE_VISITS, visits)
    ctx.set_option(capi.OPT_TILE_MIN_UNITS, 0)
    ctx.set_option(capi.OPT_SUPER_TILES, super_tiles)
    for k, (q, t) in enumerate(poses):
        ctx.set_sensor(elev, b, e, H, q, t, index=k)
    A = [IDENT.copy() for _ in meshes]
    for gi, (name, v, tr) in enumerate(meshes):
        add_mesh(ctx, name, v, tr, A[gi])
    for it in range(2):
        assert ctx.commit_scene(capi.COMMIT_AUTO) == 0
        s0 = lso.Sensor(elev, b, e, H, poses[0][0], poses[0][1])
        verts, tris = lso.concat_scene(s0, [(v, tr, A[gi]) for gi, (_, v, tr) in enumerate(meshes)], mode=1)
        bvh = lso.Bvh(verts, tris)
        ctx.set_option(capi.OPT_TRACE_STATS, 1)
        assert ctx.trace_batch(0, len(poses), capi.TRACE_NO_POINTS) == 0
        cnt = ctx.trace_counters()
        assert cnt["tile_entries"] + cnt["tile_fallbacks"] > 0 and (entries < 8 or cnt["tile_entries"] > 0)
        ctx.set_option(capi.OPT_TRACE_STATS, 0)
        assert ctx.trace_batch(0, len(poses), capi.TRACE_ORDERED) == 0
        for k, (q, t) in enumerate(poses):
            s = lso.Sensor(elev, b, e, H, q, t)
            dirs_s = lso.ray_dirs(s.tables)
            t_ref, tri_ref = bvh.trace(s.T, lso.rotate_dirs(s.R, dirs_s))
            rc, pts = ctx.fetch_points(k, capi.TRACE_ORDERED)
            assert rc == 0 and np.array_equal(pts, lso.pack_points(dirs_s, t_ref, tri_ref, H))
            rc, tt, tri = ctx.trace_raw(k)
            assert rc == 0
            assert np.array_equal(tri, tri_ref) and np.array_equal(tt, t_ref)
        if it == 1:
            # detection radius 60 m + azimuth window on sensor 1: same as filtering the unlimited result
            s = lso.Sensor(elev, b, e, H, poses[1][0], poses[1][1])
            t_ref, tri_ref = bvh.trace(s.T, lso.rotate_dirs(s.R, lso.ray_dirs(s.tables)))
            assert ctx.set_sensor_limits(60.0, 30.0, 170.0, index=1) == 0
            rc, tt, tri = ctx.trace_raw(1)
            assert rc == 0
            t_lim, tri_lim = lso.apply_limits(s, t_ref, tri_ref, 60.0, (30.0, 170.0))
            assert np.array_equal(tri, tri_lim) and np.array_equal(tt, t_lim)
        bvh.close()
        for gi in range(1, len(meshes)):
            A[gi] = lso.affine_from_components((0.3 * gi, -0.2 * gi, 0.0), (0.0, 0.0, 0.01 * gi))
            assert ctx.update_geometry(meshes[gi][0], None, None, A[gi]) == 0
    ctx.close()


def test_tile_lists_follow_queue_rebuilds():
    """Tile intervals and tile lists are indexed by queue slot, and the queues are rebuilt (in an order that atomics
    decide) whenever another sensor range is traced: switching to a variant without lists, tracing other ranges and
    coming back must not reuse lists of the old queue order."""
    meshes = synth.city(n_boxes=380, ground_cells=200, extent=20 * synth.CITY_PITCH, n_groups=4, lots=20)
    poses = synth.sensor_lattice(2, 10 * synth.CITY_PITCH)
    elev, b, e, H = synth.lidar_table(64, 1801)
    ctx = capi.Context(0, capi.FRAME_WORLD, lab=True)
    ctx.set_option(capi.OPT_TILE_MIN_UNITS, 0)
    for k, (q, t) in enumerate(poses):
        ctx.set_sensor(elev, b, e, H, q, t, index=k)
    for name, v, tr in meshes:
        add_mesh(ctx, name, v, tr)
    assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0
    ctx.set_option(capi.OPT_TRAVERSAL, 1)
    ref = [ctx.trace_raw(k)[1:] for k in range(len(poses))]
    for rep in range(3):
        ctx.set_option(capi.OPT_TRAVERSAL, 6)
        for k in (0, 2):
            rc, t, tri = ctx.trace_raw(k)
            assert rc == 0 and np.array_equal(t, ref[k][0]) and np.array_equal(tri, ref[k][1])
        ctx.set_option(capi.OPT_TRAVERSAL, 0)              # same queues, no lists: rebuilds them for other ranges
        for k in (1, 2, 3, 2):
            rc, t, tri = ctx.trace_raw(k)
            assert rc == 0 and np.array_equal(t, ref[k][0]) and np.array_equal(tri, ref[k][1])
        ctx.set_option(capi.OPT_TRAVERSAL, 6)              # range (2, 1) again: queue cache hit, lists must be rebuilt
        rc, t, tri = ctx.trace_raw(2)
        assert rc == 0 and np.array_equal(t, ref[2][0]) and np.array_equal(tri, ref[2][1])
    ctx.close()


def test_city_10m_size_independent_properties():
    """BASELINE config 4 at full size (10 M triangles): properties that need no CPU re-trace.
    (1) every variant returns the same (t, triangle) bits; (2) each hit point t*d lies on its triangle (float64
    barycentric check) and no vertex of the hit triangle is behind the origin plane; (3) refit == rebuild."""
    meshes = synth.city()
    poses = synth.sensor_lattice(8)
    elev, b, e, H = synth.lidar_table()
    ctx = capi.Context(0, capi.FRAME_WORLD, lab=True)
    pick = (0, 27, 63)
    for k, si in enumerate(pick):
        ctx.set_sensor(elev, b, e, H, poses[si][0], poses[si][1], index=k)
    for name, v, tr in meshes:
        add_mesh(ctx, name, v, tr)
    assert sum(m[2].shape[0] for m in meshes) == 10_000_000
    assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0
    res = {}
    for mode in (0, 1, 4, 5, 6, 7, 2):
        ctx.set_option(capi.OPT_TRAVERSAL, mode)
        ctx.set_option(capi.OPT_TILE_WIDTH, 4 if mode == 1 else 8)
        res[mode] = [ctx.trace_raw(k)[1:] for k in range(len(pick))]
    for mode in (1, 4, 5, 6, 7, 2):
        for k in range(len(pick)):
            assert np.array_equal(res[mode][k][0], res[0][k][0]) and np.array_equal(res[mode][k][1], res[0][k][1])
    # variant 6 with the per-tile entry lists off and at other list sizes
    ctx.set_option(capi.OPT_TRAVERSAL, 6)
    for entries, sup in ((0, 1), (4, 1), (10, 0), (16, 1)):
        ctx.set_option(capi.OPT_TILE_ENTRIES, entries)
        ctx.set_option(capi.OPT_SUPER_TILES, sup)
        for k in range(len(pick)):
            rc, t, tri = ctx.trace_raw(k)
            assert np.array_equal(t, res[0][k][0]) and np.array_equal(tri, res[0][k][1])
    ctx.set_option(capi.OPT_TILE_ENTRIES, 10)
    assert ctx.commit_scene(capi.COMMIT_REFIT) == 0 and ctx.stats().last_commit_was_rebuild == 0
    for k in range(len(pick)):
        rc, t, tri = ctx.trace_raw(k)
        assert np.array_equal(t, res[0][k][0]) and np.array_equal(tri, res[0][k][1])
    # geometric check of the hits of one sensor
    V = np.concatenate([m[1] for m in meshes]).astype(np.float64)
    off = np.cumsum([0] + [m[1].shape[0] for m in meshes])[:-1]
    Tg = np.concatenate([m[2].astype(np.int64) + o for m, o in zip(meshes, off)])
    t, tri = res[0][1]
    s = lso.Sensor(elev, b, e, H, poses[pick[1]][0], poses[pick[1]][1])
    d = lso.rotate_dirs(s.R, lso.ray_dirs(s.tables)).astype(np.float64)
    hit = tri >= 0
    assert hit.mean() > 0.9
    P = np.asarray(s.T, np.float64)[None, :] + t[hit, None].astype(np.float64) * d[hit]
    a, bb, c = V[Tg[tri[hit], 0]], V[Tg[tri[hit], 1]], V[Tg[tri[hit], 2]]
    n = np.cross(bb - a, c - a)
    area2 = np.linalg.norm(n, axis=1)
    dist = np.abs(np.einsum("ij,ij->i", P - a, n)) / area2          # distance to the triangle's plane
    assert dist.max() < 2e-3, dist.max()
    # barycentric coordinates inside [-tol, 1+tol]
    def bary(p, q, r):
        return np.einsum("ij,ij->i", np.cross(q - p, r - p), n) / (area2 * area2)
    w0, w1, w2 = bary(P, bb, c), bary(P, c, a), bary(P, a, bb)
    tol = 2e-3
    assert min(w0.min(), w1.min(), w2.min()) > -tol and abs((w0 + w1 + w2) - 1).max() < 1e-6
    ctx.close()


def test_city_10m_vs_oracle_and_float64_second_opinion():
    """BASELINE config 4 at FULL size -- the benchmarked configuration: 10 M triangles, world frame, product library.
    Three of the 64 sensors (a corner, the middle, the opposite corner) are compared ray by ray with the oracle's CPU
    BVH (hit mask, triangle id AND t bit-exact), under a non-identity per-geometry transform (the bench's frame
    motion), and the fp32 answer is then held against the float64 second opinion: rays that differ are the grazing
    rays north_star wants counted -- each must lie within k ulp of an edge / be a tie / be seen edge-on."""
    from oracle import grazing
    from bench import frame_affine
    meshes = synth.city()
    poses = synth.sensor_lattice(8)
    elev, b, e, H = synth.lidar_table()
    pick = (0, 27, 63)
    ctx = capi.Context(0, capi.FRAME_WORLD)
    for k, si in enumerate(pick):
        ctx.set_sensor(elev, b, e, H, poses[si][0], poses[si][1], index=k)
    A = [frame_affine(3, gi) for gi in range(len(meshes))]
    for (name, v, tr), a in zip(meshes, A):
        add_mesh(ctx, name, v, tr, a)
    assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0
    s0 = lso.Sensor(elev, b, e, H, poses[0][0], poses[0][1])
    verts, tris = lso.concat_scene(s0, [(v, t, a) for (_, v, t), a in zip(meshes, A)], mode=1)
    assert tris.shape[0] == 10_000_000
    bvh = lso.Bvh(verts, tris)
    total_grazing = 0
    for k, si in enumerate(pick):
        s = lso.Sensor(elev, b, e, H, poses[si][0], poses[si][1])
        dirs = lso.rotate_dirs(s.R, lso.ray_dirs(s.tables))
        t_ref, tri_ref = bvh.trace(s.T, dirs)
        rc, t, tri = ctx.trace_raw(k)
        assert rc == 0
        assert int((tri != tri_ref).sum()) == 0, f"sensor {si}: {int((tri != tri_ref).sum())} rays differ in triangle id"
        assert np.array_equal(t, t_ref), f"sensor {si}: t differs"
        rep = grazing.second_opinion(verts, tris, s.T, dirs, t, tri, bvh=bvh)
        assert rep["unexplained"] == 0, rep
        assert rep["max_rel_range_error_vs_f64"] < 1e-4, rep              # north_star's range tolerance, vs exact arithmetic
        total_grazing += rep["grazing_rays"]
    assert total_grazing < 1000                                            # a handful per 460 928 rays
    bvh.close()
    ctx.close()


def test_heightfield_1m_float64_second_opinion():
    """BASELINE config 3: the GPU's fp32 answer against the float64 second opinion.  The sensor stands above a grid
    vertex, so the azimuth 0 / 90 / 180 / 270 rays run exactly along shared edges: the grazing report must find them
    and explain every one (edge / tie)."""
    from oracle import grazing
    v, tr = synth.heightfield()
    elev, b, e, H = synth.lidar_table()
    T = (0.0, 0.0, 2.0)
    ctx = capi.Context(0, capi.FRAME_WORLD)
    ctx.set_sensor(elev, b, e, H, (1, 0, 0, 0), T)
    add_mesh(ctx, "terrain", v, tr)
    assert ctx.commit_scene() == 0
    rc, t, tri = ctx.trace_raw()
    assert rc == 0
    s = lso.Sensor(elev, b, e, H, (1, 0, 0, 0), T)
    verts, tris = lso.concat_scene(s, [(v, tr, IDENT)], mode=1)
    dirs = lso.rotate_dirs(s.R, lso.ray_dirs(s.tables))
    bvh = lso.Bvh(verts, tris)
    t_ref, tri_ref = bvh.trace(s.T, dirs)
    assert np.array_equal(tri, tri_ref) and np.array_equal(t, t_ref)
    rep = grazing.second_opinion(verts, tris, s.T, dirs, t, tri, bvh=bvh)
    bvh.close()
    assert rep["unexplained"] == 0, rep
    assert 0 < rep["grazing_rays"] < 500, rep
    assert rep["max_rel_range_error_vs_f64"] < 1e-4, rep
    ctx.close()


def test_trajectory_batch_moving_meshes_rebuild_each_frame(meshes):
    """BASELINE config 5 at reduced frame count: the 1M-triangle heightfield + 16 instances of `ben`,
    each following the looped trajectory.json twists with a phase offset of 0.9 s, FULL LBVH rebuild every frame,
    128 x 3601 rays, frames sharded round-robin over 2 ranks (this process plays rank 1).  Every traced frame is
    compared bit-for-bit (t, triangle id, ordered points) with the oracle's CPU BVH on the same poses."""
    from lidarshooter_b200 import shard
    hv, ht = synth.heightfield()
    bv, bt = meshes["ben_v"], meshes["ben_t"]
    loop = synth.trajectory_poses([(10.0, (0, 5, 0), (0, 0, 1)), (5.0, (1, 0, 0), (0, 0, -1))])     # 150 ticks
    n_inst, n_frames_total = 16, 12
    mine = [f for d, f in shard.shard_scans(1, n_frames_total, rank=1, world=2, major="frame")]
    assert mine == [1, 3, 5, 7, 9, 11]
    elev, b, e, H = synth.lidar_table()
    T = (0.0, 0.0, 4.0)
    ctx = capi.Context(0, capi.FRAME_WORLD)
    ctx.set_sensor(elev, b, e, H, (1, 0, 0, 0), T)
    s = lso.Sensor(elev, b, e, H, (1, 0, 0, 0), T)
    add_mesh(ctx, "terrain", hv, ht)
    base = [np.array([(i % 4 - 1.5) * 12.0, (i // 4 - 1.5) * 12.0, 3.0], np.float32) for i in range(n_inst)]
    for i in range(n_inst):
        add_mesh(ctx, f"ben{i:02d}", bv, bt)

    def affines(frame):
        A = np.zeros((1 + n_inst, 12), np.float32)
        A[0] = IDENT
        for i in range(n_inst):
            lin, ang = loop[(frame + 9 * i) % len(loop)]
            A[1 + i] = lso.affine_from_components(lin * np.float32(0.2) + base[i], ang)
        return A

    A = np.stack([affines(f) for f in mine])
    R = ctx.ray_count(0)
    outs = [[np.zeros((R, 32), np.uint8)] for _ in mine]
    rc, counts = ctx.run_frames(A, [[o[0].ctypes.data] for o in outs], [R], first=0, count=1,
                                commit_mode=capi.COMMIT_REBUILD, flags=capi.TRACE_ORDERED)
    assert rc == 0, ctx.last_error()
    assert ctx.stats().last_commit_was_rebuild == 1
    assert int(ctx.stats().n_triangles) == 1_000_000 + n_inst * bt.shape[0]
    for k, f in enumerate(mine[:3] + mine[-1:]):
        kk = mine.index(f)
        ref = lso.scan(s, [(hv, ht, IDENT)] + [(bv, bt, A[kk, 1 + i]) for i in range(n_inst)], mode=1, use_bvh=True)
        assert counts[kk, 0] == ref["points"].shape[0]
        assert np.array_equal(outs[kk][0][:counts[kk, 0]], ref["points"])
    # the last frame is still resident: per-ray results too
    rc, t, tri = ctx.trace_raw(0)
    assert np.array_equal(t, ref["t"]) and np.array_equal(tri, ref["tri"])
    assert (tri >= 1_000_000).sum() > 100          # some rays do hit the moving meshes
    ctx.close()


@pytest.mark.parametrize("mode", [6, 1, 2])
def test_sensor_limits_detection_radius_and_azimuth_window(meshes, mode):
    """detection radius and horizontal angle range (config keys the reference never applies), as
    lsb200_set_sensor_limits defines them; bit-exact against the oracle with the same limits; off by default."""
    args = xt32_sensor_args("0000")
    s = lso.Sensor(*args)
    ms = [(meshes["ground_v"], meshes["ground_t"], IDENT), (meshes["ben_v"], meshes["ben_t"], IDENT)]
    ctx = make_ctx("0000", opts={capi.OPT_TRAVERSAL: mode}, lab=(mode != 6))
    add_mesh(ctx, "ground", meshes["ground_v"], meshes["ground_t"])
    add_mesh(ctx, "ben", meshes["ben_v"], meshes["ben_t"])
    assert ctx.commit_scene() == 0
    base = lso.scan(s, ms)
    r_med = float(np.median(base["t"][base["tri"] >= 0]))
    for max_range, win in ((0.0, (-180.0, 180.0)), (r_med, (-180.0, 180.0)), (0.0, (-90.0, 45.5)), (1.5 * r_med, (100.0, 180.0)),
                           (1000.0, (-180.0, 180.0))):
        assert ctx.set_sensor_limits(max_range, win[0], win[1]) == 0
        rc, pts = ctx.trace_scene(0, capi.TRACE_ORDERED)
        rc2, t, tri = ctx.trace_raw()
        ref = lso.scan(s, ms, max_range=max_range, az_window=win)
        assert rc == 0 and rc2 == 0
        assert np.array_equal(t, ref["t"]) and np.array_equal(tri, ref["tri"]) and np.array_equal(pts, ref["points"])
        if max_range == r_med:
            assert 0 < pts.shape[0] < base["points"].shape[0] and float(t[tri >= 0].max()) <= np.float32(r_med)
        if win == (-90.0, 45.5):
            assert 0 < pts.shape[0] < base["points"].shape[0]
        if max_range in (0.0, 1000.0) and win == (-180.0, 180.0):
            assert pts.shape[0] == 1781
    assert ctx.set_sensor_limits(5.0, 10.0, -10.0) == capi.ERR_INVALID
    # set_sensor clears the limits
    ctx.set_sensor(*args)
    assert ctx.commit_scene() == 0
    assert ctx.trace_scene(0, capi.TRACE_ORDERED)[1].shape[0] == 1781
    ctx.close()


@pytest.mark.parametrize("n_tris", [1, 2, 255, 256, 257, 20000, 300001])
def test_fit_by_range_queries_equals_climb(n_tris):
    """k_fit_rmq (block-local node boxes from a sparse table over the block's leaf boxes) writes the same node array as
    the bottom-up climb of k_fit, for rebuild and refit (multi-kernel pipeline forced)."""
    rng = np.random.default_rng(n_tris)
    c = rng.uniform(-50, 50, (n_tris, 1, 3)).astype(np.float32)
    c[: n_tris // 4] = c[0]                                   # long runs of equal Morton codes
    v = (c + rng.normal(0, 0.4, (n_tris, 3, 3)).astype(np.float32)).reshape(-1, 3)
    tr = np.arange(3 * n_tris, dtype=np.uint32).reshape(-1, 3)
    A2 = lso.affine_from_components([1.5, -0.5, 0.25], [0.2, 0.1, -0.3])
    out = []
    for rmq in (1, 0):
        ctx = capi.Context(0, lab=True)
        ctx.set_option(capi.OPT_SMALL_BUILD, 0)
        ctx.set_option(capi.OPT_FUSED_TREE, 0)
        ctx.set_option(capi.OPT_FIT_RMQ, rmq)
        ctx.set_option(capi.OPT_LEAF_SIZE, 4)
        ctx.set_sensor(np.linspace(40, -40, 16).astype(np.float32), 0.0, 360.0, 64, (1, 0, 0, 0), (0, 0, 0))
        add_mesh(ctx, "soup", v, tr)
        assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0, ctx.last_error()
        nodes, ids, root = ctx.debug_copy_bvh()
        rc, t, tri = ctx.trace_raw()
        assert ctx.update_transform("soup", A2) == 0 and ctx.commit_scene(capi.COMMIT_REFIT) == 0
        nodes2, _, _ = ctx.debug_copy_bvh()
        rc, t2, tri2 = ctx.trace_raw()
        out.append((nodes.copy(), ids.copy(), t, tri, nodes2.copy(), t2, tri2))
        ctx.close()
    for name, x, y in zip(("nodes", "leaf ids", "t", "tri", "nodes after refit", "t after refit", "tri after refit"), *out):
        assert np.array_equal(np.atleast_1d(x).view(np.uint8), np.atleast_1d(y).view(np.uint8)), name
