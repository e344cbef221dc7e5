"""GPU tests of the round-2 boundary additions, all through the C ABI (include/lsb200.h) and checked against oracle/:
log sink, several contexts / devices in one process, lsb200_clone_scene + lsb200_run_batch, range-image output and
its host expander, pinned host memory from the library, watertight mode (Woop 2013) with a leak count, intensity
model + minimum-intensity filter, empty-scene device results."""
import numpy as np
import pytest

from lidarshooter_b200 import capi, synth
from oracle import lso
from conftest import xt32_sensor_args

pytestmark = pytest.mark.gpu
IDENT = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)


def n_gpus():
    import torch
    return torch.cuda.device_count()


def xt32_ctx(meshes, uid="0000", device=0, opts=None):
    ctx = capi.Context(device)
    for k, v in (opts or {}).items():
        ctx.set_option(k, v)
    ctx.set_sensor(*xt32_sensor_args(uid))
    for name in ("ground", "ben"):
        v, t = meshes[name + "_v"], meshes[name + "_t"]
        assert ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0]) >= 0
        assert ctx.update_geometry(name, v, t, IDENT) == 0
    return ctx


def xt32_meshes(meshes):
    return [(meshes["ground_v"], meshes["ground_t"], IDENT), (meshes["ben_v"], meshes["ben_t"], IDENT)]


# ---- several contexts / devices in one process (VERDICT weak #6, ADVICE medium) -------------------------------------
def test_contexts_on_every_device_of_the_process():
    """One process, one context per visible GPU (twice device 0 when there is only one), each with a scene large
    enough to need the > 48 KB dynamic-shared-memory kernels (onesweep sort; a 16 000-triangle fused small build):
    the per-device function attributes must have been set for EVERY device, and the results must agree."""
    devices = list(range(n_gpus())) if n_gpus() > 1 else [0, 0]
    rng = np.random.default_rng(3)
    big_v, big_t = synth.heightfield(nx=200, ny=150)            # 60 000 triangles: multi-kernel pipeline, onesweep
    small_v = rng.uniform(-20, 20, (48_000, 3)).astype(np.float32)
    small_t = np.arange(48_000, dtype=np.uint32).reshape(-1, 3)  # 16 000 triangles: k_build_small with 64 000 B of smem
    elev, b, e, H = synth.lidar_table(32, 361)
    results = []
    ctxs = []
    for dev in devices:
        per_dev = []
        for v, t in ((big_v, big_t), (small_v, small_t)):
            ctx = capi.Context(dev)
            ctxs.append(ctx)
            ctx.set_sensor(elev, b, e, H, (1, 0, 0, 0), (0.0, 0.0, 3.0))
            assert ctx.add_geometry("m", capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0]) >= 0
            assert ctx.update_geometry("m", v, t, IDENT) == 0
            assert ctx.commit_scene() == 0, ctx.last_error()
            rc, tt, tri = ctx.trace_raw()
            assert rc == 0, ctx.last_error()
            assert (tri >= 0).sum() > 100
            per_dev.append((tt, tri))
        results.append(per_dev)
    for per_dev in results[1:]:
        for (t0, i0), (t1, i1) in zip(results[0], per_dev):
            assert np.array_equal(t0, t1) and np.array_equal(i0, i1)
    for c in ctxs:
        c.close()


# ---- log sink ------------------------------------------------------------------------------------------------------
def test_log_callback_receives_errors_and_lifecycle(meshes):
    ctx = xt32_ctx(meshes)
    lines = []
    assert ctx.set_log_callback(lambda level, msg: lines.append((level, msg))) == 0
    assert ctx.commit_scene() == 0
    assert any(lv == capi.LOG_DEBUG and "commit" in m and "5651" in m for lv, m in lines), lines
    assert ctx.remove_geometry("no-such-mesh") == -1
    assert lines[-1][0] == capi.LOG_ERROR and lines[-1][1] == ctx.last_error() and "no-such-mesh" in lines[-1][1]
    n = len(lines)
    ctx.set_log_callback(None)
    assert ctx.remove_geometry("no-such-mesh") == -1
    assert len(lines) == n
    ctx.close()


# ---- range image ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("flags", [0, capi.TRACE_XYZI16, capi.TRACE_RING_ZERO, capi.TRACE_WORLD_POINTS | capi.TRACE_XYZI16])
def test_range_image_and_host_expander(meshes, flags):
    """LSB200_TRACE_RANGE_IMAGE delivers the per-ray t array (4 B / ray); lsb200_expand_range_image rebuilds, on the
    host, exactly the bytes an LSB200_TRACE_ORDERED trace writes on the device."""
    ctx = xt32_ctx(meshes)
    assert ctx.commit_scene() == 0
    rc, t_raw, tri = ctx.trace_raw()
    rc, img = ctx.trace_scene(7, capi.TRACE_RANGE_IMAGE)
    assert rc == 0 and img.shape == (4800, 4)
    t_img = img.reshape(-1).view(np.float32)
    assert np.array_equal(t_img.view(np.uint32), t_raw.view(np.uint32))
    ref = lso.scan(lso.Sensor(*xt32_sensor_args("0000")), xt32_meshes(meshes))
    assert np.array_equal(t_img, ref["t"])
    rc, dev_pts = ctx.trace_scene(7, capi.TRACE_ORDERED | flags)
    rc2, host_pts = ctx.expand_range_image(t_img, flags)
    assert rc == 0 and rc2 == 0 and dev_pts.shape[0] == 1781
    assert np.array_equal(dev_pts, host_pts)
    ctx.close()


def test_run_frames_range_image_matches_per_frame_calls(meshes):
    """The pipelined batch entry with range-image output (two device slots, copies overlapping the next frame)."""
    import torch
    poses = synth.trajectory_poses([(0.5, (0, 5, 0), (0, 0, 1))])
    ctx = xt32_ctx(meshes)
    R = ctx.ray_count()
    A = np.stack([np.stack([IDENT, capi.affine_from_components(tr, eu)]) for tr, eu in poses])
    outs = [torch.empty(R, dtype=torch.float32).pin_memory() for _ in poses]
    rc, cnt = ctx.run_frames(A, [[o.data_ptr()] for o in outs], [R], first=0, count=1, commit_mode=capi.COMMIT_AUTO,
                             flags=capi.TRACE_RANGE_IMAGE)
    assert rc == 0, ctx.last_error()
    assert (cnt == R).all()
    s = lso.Sensor(*xt32_sensor_args("0000"))
    for f, (tr, eu) in enumerate(poses):
        ref = lso.scan(s, [(meshes["ground_v"], meshes["ground_t"], IDENT),
                           (meshes["ben_v"], meshes["ben_t"], lso.affine_from_components(tr, eu))])
        assert np.array_equal(outs[f].numpy(), ref["t"]), f
    ctx.close()


def test_alloc_host_is_usable_as_output(meshes):
    ctx = xt32_ctx(meshes)
    assert ctx.commit_scene() == 0
    buf = ctx.alloc_host(4800 * 32)
    assert ctx.numa_node() >= -1
    rc, pts = ctx.trace_scene(0, capi.TRACE_ORDERED, out=buf.reshape(4800, 32))
    assert rc == 0 and pts.shape[0] == 1781
    ref = lso.scan(lso.Sensor(*xt32_sensor_args("0000")), xt32_meshes(meshes))
    assert np.array_equal(pts, ref["points"])
    ctx.free_host(buf)
    ctx.close()


# ---- clone + multi-GPU batch ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("sharding", [capi.SHARD_FRAMES, capi.SHARD_SENSORS])
def test_clone_scene_and_run_batch(sharding):
    """lsb200_clone_scene replicates the scene on another GPU (device 1 when present, else a second context on device
    0); lsb200_run_batch over both contexts returns, frame by frame and sensor by sensor, the clouds a single context's
    lsb200_run_frames returns."""
    import torch
    meshes = synth.city(n_boxes=120, ground_cells=80, extent=12 * synth.CITY_PITCH, n_groups=3, lots=12)
    poses = synth.sensor_lattice(2, 4 * synth.CITY_PITCH)
    elev, b, e, H = synth.lidar_table(16, 181)
    ctx = capi.Context(0, capi.FRAME_WORLD)
    for k, (q, t) in enumerate(poses):
        ctx.set_sensor(elev, b, e, H, q, t, index=k)
    ctx.set_sensor_limits(150.0, -170.0, 170.0, index=1)
    for name, v, tr in meshes:
        assert ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], tr.shape[0]) >= 0
        assert ctx.update_geometry(name, v, tr, IDENT) == 0
    other = ctx.clone(1 if n_gpus() > 1 else 0)
    assert other.geometry_count() == ctx.geometry_count() and other.ray_count(3) == ctx.ray_count(3)
    F, S, R = 5, len(poses), elev.size * H
    A = np.zeros((F, len(meshes), 12), np.float32)
    for f in range(F):
        for gi in range(len(meshes)):
            A[f, gi] = IDENT
            if gi:
                A[f, gi, 3] = 0.3 * f + 0.1 * gi
    flags = capi.TRACE_ORDERED | capi.TRACE_XYZI16

    def buffers():
        bufs = [[torch.zeros((R, 16), dtype=torch.uint8).pin_memory() for _ in range(S)] for _ in range(F)]
        return bufs, [[t.data_ptr() for t in row] for row in bufs]

    b1, p1 = buffers()
    rc, c1 = ctx.run_frames(A, p1, [R] * S, first=0, count=S, commit_mode=capi.COMMIT_REBUILD, flags=flags)
    assert rc == 0, ctx.last_error()
    b2, p2 = buffers()
    rc, c2 = capi.run_batch([ctx, other], A, p2, [R] * S, first=0, count=S, commit_mode=capi.COMMIT_REBUILD, flags=flags,
                            sharding=sharding)
    assert rc == 0, (ctx.last_error(), other.last_error())
    assert np.array_equal(c1, c2) and c1.sum() > 1000
    for f in range(F):
        for k in range(S):
            n = int(c1[f, k])
            assert np.array_equal(b1[f][k].numpy()[:n], b2[f][k].numpy()[:n]), (f, k)
    other.close()
    ctx.close()


# ---- watertight mode (VERDICT missing #1 / N1) ---------------------------------------------------------------------
@pytest.mark.parametrize("uid", ["0000", "0001"])
def test_watertight_mode_xt32_vs_oracle(meshes, uid):
    """LSB200_OPT_WATERTIGHT: GPU == oracle in the same mode, bit for bit (mask, id, t), for the traversal and for brute
    force; against the default Moeller-Trumbore mode the hit set of the reference scene is the same and ranges agree to
    1e-5 (both are correct intersectors; they differ only on grazing rays)."""
    s = lso.Sensor(*xt32_sensor_args(uid))
    ref_mt = lso.scan(s, xt32_meshes(meshes))
    with lso.watertight():
        ref_wt = lso.scan(s, xt32_meshes(meshes))
    for mode in (6, 3):
        ctx = xt32_ctx(meshes, uid, opts={capi.OPT_WATERTIGHT: 1, capi.OPT_TRAVERSAL: mode})
        assert ctx.commit_scene() == 0
        rc, t, tri = ctx.trace_raw()
        assert rc == 0, ctx.last_error()
        assert np.array_equal(tri, ref_wt["tri"]) and np.array_equal(t, ref_wt["t"])
        rc, pts = ctx.trace_scene(0, capi.TRACE_ORDERED)
        assert np.array_equal(pts, ref_wt["points"])
        # back to the default intersector on the same context: records are rewritten by the next commit
        ctx.set_option(capi.OPT_WATERTIGHT, 0)
        assert ctx.commit_scene() == 0
        rc, t0, tri0 = ctx.trace_raw()
        assert np.array_equal(tri0, ref_mt["tri"]) and np.array_equal(t0, ref_mt["t"])
        ctx.close()
    assert np.array_equal(ref_wt["tri"] >= 0, ref_mt["tri"] >= 0)
    hit = ref_mt["tri"] >= 0
    assert np.abs(ref_wt["t"][hit] - ref_mt["t"][hit]).max() / ref_mt["t"][hit].max() < 1e-5


def ray_aligned_shell(sensor, radius=20.0):
    """A mesh whose every vertex lies (to fp32 rounding) ON one ray of the sensor: vertex (v, h) = radius * d(v, h),
    the (v, h) grid triangulated.  Every ray then grazes six triangles at their common vertex -- the worst case for
    a non-watertight intersector."""
    d = lso.ray_dirs(sensor.tables).reshape(sensor.V, sensor.H, 3)
    r = (radius * (1.0 + 0.2 * np.sin(np.arange(sensor.V))[:, None] * np.cos(0.3 * np.arange(sensor.H))[None, :])).astype(np.float32)
    verts = (r[:, :, None] * d).astype(np.float32).reshape(-1, 3)
    Hh = sensor.H - 1                                       # the last azimuth step repeats the first direction
    v, h = np.meshgrid(np.arange(sensor.V - 1), np.arange(Hh - 1), indexing="ij")
    v00 = (v * sensor.H + h).astype(np.uint32)
    tris = np.concatenate([np.stack([v00, v00 + 1, v00 + sensor.H + 1], -1).reshape(-1, 3),
                           np.stack([v00, v00 + sensor.H + 1, v00 + sensor.H], -1).reshape(-1, 3)]).astype(np.uint32)
    interior = np.zeros((sensor.V, sensor.H), bool)
    interior[1:-1, 1:Hh - 1] = True
    return verts, tris, interior.reshape(-1)


def test_watertight_mode_closes_the_leaks_of_moeller_trumbore():
    """The leak count north_star asks to be reported: rays that MISS in the reference's Moeller-Trumbore mode although
    they pass through the interior of a closed patch of triangles, and HIT in watertight mode.  Scene: the ray-aligned
    shell (every interior ray pierces a vertex shared by six triangles) around an XT-32 and around the 128 x 3601 device.
    GPU == oracle bit-exactly in both modes; watertight mode has zero leaks."""
    for table, tf in ((xt32_sensor_args("0000")[:4], ((1, 0, 0, 0), (0, 0, 0))), (synth.lidar_table(), ((1, 0, 0, 0), (0, 0, 0)))):
        s = lso.Sensor(*table, *tf)
        verts, tris, interior = ray_aligned_shell(s)
        leaks = {}
        for wt in (0, 1):
            ctx = capi.Context(0)
            ctx.set_option(capi.OPT_WATERTIGHT, wt)
            ctx.set_sensor(*table, *tf)
            assert ctx.add_geometry("shell", capi.GEOMETRY_TRIANGLE, verts.shape[0], tris.shape[0]) >= 0
            assert ctx.update_geometry("shell", verts, tris, IDENT) == 0
            assert ctx.commit_scene() == 0
            rc, t, tri = ctx.trace_raw()
            assert rc == 0
            with lso.watertight(bool(wt)):
                ref = lso.scan(s, [(verts, tris, IDENT)], use_bvh=True)
            assert np.array_equal(tri, ref["tri"]) and np.array_equal(t, ref["t"])
            leaks[wt] = int(((tri < 0) & interior).sum())
            ctx.close()
        print(f"leak count, {s.V} x {s.H} rays through shared vertices: Moeller-Trumbore {leaks[0]}, watertight {leaks[1]}")
        assert leaks[1] == 0
        assert leaks[0] > 0            # the reference's intersector does leak on this scene


def test_watertight_heightfield_leak_count():
    """Config 3 (1 M-triangle heightfield, 460 928 rays): rays that miss in Moeller-Trumbore mode but hit in watertight
    mode = leaks of the reference's intersector on a benchmark scene (reported; few or none -- the LiDAR's directions
    are not aimed at edges), and the two modes otherwise agree: same triangle wherever both hit away from an edge."""
    v, tr = synth.heightfield()
    elev, b, e, H = synth.lidar_table()
    res = {}
    for wt in (0, 1):
        ctx = capi.Context(0, capi.FRAME_WORLD)
        ctx.set_option(capi.OPT_WATERTIGHT, wt)
        ctx.set_sensor(elev, b, e, H, (1, 0, 0, 0), (0.0, 0.0, 2.0))
        assert ctx.add_geometry("terrain", capi.GEOMETRY_TRIANGLE, v.shape[0], tr.shape[0]) >= 0
        assert ctx.update_geometry("terrain", v, tr, IDENT) == 0
        assert ctx.commit_scene() == 0
        rc, t, tri = ctx.trace_raw()
        assert rc == 0
        res[wt] = (t, tri)
        ctx.close()
    s = lso.Sensor(elev, b, e, H, (1, 0, 0, 0), (0.0, 0.0, 2.0))
    verts, tris = lso.concat_scene(s, [(v, tr, IDENT)], mode=1)
    dirs = lso.rotate_dirs(s.R, lso.ray_dirs(s.tables))
    bvh = lso.Bvh(verts, tris)
    with lso.watertight():
        t_ref, tri_ref = bvh.trace(s.T, dirs)
    bvh.close()
    assert np.array_equal(res[1][1], tri_ref) and np.array_equal(res[1][0], t_ref)
    leaks = int(((res[0][1] < 0) & (res[1][1] >= 0)).sum())
    extra = int(((res[0][1] >= 0) & (res[1][1] < 0)).sum())
    differ = int((res[0][1] != res[1][1]).sum())
    print(f"heightfield: {leaks} rays leak through Moeller-Trumbore, {extra} hit only there, {differ} differ in triangle id")
    assert leaks < 100 and extra < 100 and differ < 2000


# ---- intensity model + filter --------------------------------------------------------------------
def test_intensity_model_and_minimum_intensity_filter(meshes):
    s = lso.Sensor(*xt32_sensor_args("0000"))
    base = lso.scan(s, xt32_meshes(meshes))
    dirs = lso.ray_dirs(s.tables)
    # model 1 (cos incidence), no filter
    ctx = xt32_ctx(meshes, opts={capi.OPT_INTENSITY_MODEL: 1})
    assert ctx.commit_scene() == 0
    rc, pts = ctx.trace_scene(0, 0)                       # unordered requested: an intensity model implies the ordered path
    assert rc == 0 and pts.shape[0] == 1781
    t1, tri1, I1 = lso.intensity(base["verts"], base["tris"], dirs, base["t"], base["tri"], model=1)
    rc, I = ctx.fetch_intensity()
    assert rc == 0 and np.array_equal(I, I1)
    assert 0.0 < I1[base["tri"] >= 0].min() and I1.max() <= 255.0 and np.unique(I1).size > 100
    ref_pts = lso.pack_points(dirs, t1, tri1, s.H, intensity=I1)
    assert np.array_equal(pts, ref_pts)
    # the device config's filter (use_intensity_filtering / minimum_intensity), threshold chosen to bite
    thr = float(np.median(I1[base["tri"] >= 0]))
    assert ctx.set_sensor_intensity_filter(thr) == 0
    rc, pts = ctx.trace_scene(1, capi.TRACE_ORDERED)
    t2, tri2, I2 = lso.intensity(base["verts"], base["tris"], dirs, base["t"], base["tri"], model=1, min_intensity=thr)
    assert 0 < (tri2 >= 0).sum() < 1781
    assert pts.shape[0] == int((tri2 >= 0).sum())
    assert np.array_equal(pts, lso.pack_points(dirs, t2, tri2, s.H, intensity=I2))
    rc, t, tri = ctx.trace_raw()
    assert np.array_equal(tri, tri2) and np.array_equal(t, t2)
    ctx.close()
    # the reference's constant model with its config's threshold (5 < 64): nothing is dropped, bytes unchanged
    ctx = xt32_ctx(meshes)
    assert ctx.set_sensor_intensity_filter(5.0) == 0
    assert ctx.commit_scene() == 0
    rc, pts = ctx.trace_scene(0, capi.TRACE_ORDERED)
    assert np.array_equal(pts, base["points"])
    ctx.close()


def test_empty_scene_device_results_are_misses():
    """ADVICE: after an empty-scene trace the DEVICE result arrays say 'miss' (t = +inf, triangle -1), like the host copy."""
    import ctypes as C
    import torch
    ctx = capi.Context(0)
    ctx.set_sensor(*xt32_sensor_args("0000"))
    assert ctx.trace_device(0, 0) == -1
    L = capi.lib()
    dp, dc, dt, dtri = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
    assert L.lsb200_device_results(ctx._h, 0, C.byref(dp), C.byref(dc), C.byref(dt), C.byref(dtri)) == 0
    ctx.synchronize()
    t = np.empty(4800, np.float32)
    tri = np.empty(4800, np.int32)
    rt = C.CDLL("libcudart.so")
    assert rt.cudaMemcpy(C.c_void_p(t.ctypes.data), dt, C.c_size_t(4800 * 4), 2) == 0
    assert rt.cudaMemcpy(C.c_void_p(tri.ctypes.data), dtri, C.c_size_t(4800 * 4), 2) == 0
    assert np.isinf(t).all() and (tri == -1).all()
    ctx.close()


@pytest.mark.parametrize("leaf", [1, 2, 8])
@pytest.mark.parametrize("n_tris", [2, 3, 255, 256, 257, 511, 513, 20000, 300001])
def test_fused_tree_equals_pipeline(n_tris, leaf):
    """k_tree (hierarchy + boxes in one bottom-up pass, Apetrei 2014 with Karras numbering; replaces rtcCommitScene,
    SRC/EmbreeTracer.cpp:290-295) writes the SAME node array, leaf order and per-ray results as k_hierarchy + k_fit,
    rebuild and refit, with many equal Morton codes (ties broken by index) and a ragged last block."""
    rng = np.random.default_rng(11 * n_tris + leaf)
    c = rng.uniform(-30, 30, (n_tris, 1, 3)).astype(np.float32)
    c[: n_tris // 3] = c[0]                                  # a third of the centroids coincide
    c[n_tris // 3: n_tris // 2, 0, 2] = 0.0                  # and a flat sheet (long common prefixes)
    v = (c + rng.normal(0, 0.4, (n_tris, 3, 3)).astype(np.float32)).reshape(-1, 3)
    tr = np.arange(3 * n_tris, dtype=np.uint32).reshape(-1, 3)
    elev = np.linspace(50, -50, 16).astype(np.float32)
    A2 = lso.affine_from_components([0.5, -0.25, 0.125], [0.1, -0.2, 0.3])
    out = []
    for fused in (1, 0):
        ctx = capi.Context(0)
        ctx.set_option(capi.OPT_SMALL_BUILD, 0)
        ctx.set_option(capi.OPT_FUSED_TREE, fused)
        ctx.set_option(capi.OPT_LEAF_SIZE, leaf)
        ctx.set_sensor(elev, 0.0, 360.0, 64, (0.9, 0.1, -0.2, 0.3), (1.0, -2.0, 0.5))
        assert ctx.add_geometry("soup", capi.GEOMETRY_TRIANGLE, v.shape[0], tr.shape[0]) >= 0
        assert ctx.update_geometry("soup", v, tr, IDENT) == 0
        res = []
        for rep in range(2):                                 # twice: the global arrival counters are parity-based
            assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0, ctx.last_error()
            nodes, ids, root = ctx.debug_copy_bvh()
            rc, t, tri = ctx.trace_raw()
            res += [capi.reachable_nodes(nodes, root), ids.copy(), root, t, tri]
        assert ctx.update_transform("soup", A2) == 0
        assert ctx.commit_scene(capi.COMMIT_REFIT) == 0 and ctx.stats().last_commit_was_rebuild == 0
        nodes2, ids2, root2 = ctx.debug_copy_bvh()
        rc, t2, tri2 = ctx.trace_raw()
        out.append(res + [capi.reachable_nodes(nodes2, root2), ids2.copy(), t2, tri2])
        ctx.close()
    a, b = out
    for k, (x, y) in enumerate(zip(a, b)):
        assert np.array_equal(np.atleast_1d(x).view(np.uint8), np.atleast_1d(y).view(np.uint8)), k
    assert n_tris < 100 or (a[4] >= 0).sum() > 0


@pytest.mark.parametrize("refill", [0, 15, 6])
def test_lane_refill_small_city_vs_oracle_bvh(refill):
    """Ray compaction inside the traversal (run_phase_refill: free lanes take the next rays of the queue by ballot +
    rank, unit lists in a shared-memory ring; the reference's counterpart is the per-packet loop of
    SRC/EmbreeTracer.cpp:297-367): per-ray t / triangle id bit-exact against the oracle's CPU BVH for every group of
    queues, ragged tables (721 azimuth steps: tiles that fall off the table), a detection radius + azimuth window,
    and the unordered warp-aggregated append delivers exactly the ordered cloud's points."""
    meshes = synth.city(n_boxes=380, ground_cells=200, extent=20 * synth.CITY_PITCH, n_groups=4, lots=20)
    poses = synth.sensor_lattice(2, 10 * synth.CITY_PITCH)
    elev, b, e, H = synth.lidar_table(30, 721)
    ctx = capi.Context(0, capi.FRAME_WORLD, lab=True)       # the refill phases are compiled into the lab library only
    ctx.set_option(capi.OPT_TILE_MIN_UNITS, 0)
    ctx.set_option(capi.OPT_LANE_REFILL, refill)
    for k, (q, t) in enumerate(poses):
        ctx.set_sensor(elev, b, e, H, q, t, index=k)
    for gi, (name, v, tr) in enumerate(meshes):
        assert ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], tr.shape[0]) >= 0
        assert ctx.update_geometry(name, v, tr, IDENT) == 0
    assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0
    s0 = lso.Sensor(elev, b, e, H, poses[0][0], poses[0][1])
    verts, tris = lso.concat_scene(s0, [(v, tr, IDENT) for (_, v, tr) in meshes], mode=1)
    bvh = lso.Bvh(verts, tris)
    assert ctx.set_sensor_limits(60.0, 30.0, 170.0, index=1) == 0
    assert ctx.trace_batch(0, len(poses), capi.TRACE_XYZI16) == 0            # unordered append
    unordered = [ctx.fetch_points(k, capi.TRACE_XYZI16)[1].copy() for k in range(len(poses))]
    raw = [ctx.trace_raw(k) for k in range(len(poses))]
    assert ctx.trace_batch(0, len(poses), capi.TRACE_XYZI16 | capi.TRACE_ORDERED) == 0
    for k, (q, t) in enumerate(poses):
        s = lso.Sensor(elev, b, e, H, q, t)
        t_ref, tri_ref = bvh.trace(s.T, lso.rotate_dirs(s.R, lso.ray_dirs(s.tables)))
        if k == 1:
            t_ref, tri_ref = lso.apply_limits(s, t_ref, tri_ref, 60.0, (30.0, 170.0))
        rc, tt, tri = raw[k]
        assert rc == 0 and np.array_equal(tri, tri_ref) and np.array_equal(tt, t_ref), k
        rc, tt, tri = ctx.trace_raw(k)
        assert rc == 0 and np.array_equal(tri, tri_ref) and np.array_equal(tt, t_ref), k
        rc, ordered = ctx.fetch_points(k, capi.TRACE_XYZI16 | capi.TRACE_ORDERED)
        assert rc == 0 and ordered.shape[0] == int((tri_ref >= 0).sum())
        a = np.ascontiguousarray(unordered[k]).view(np.uint8).reshape(-1, 16)
        o = np.ascontiguousarray(ordered).view(np.uint8).reshape(-1, 16)
        assert a.shape == o.shape
        key = lambda x: x[np.lexsort(x.T[::-1])]
        assert np.array_equal(key(a), key(o)), k
    bvh.close()
    ctx.close()


@pytest.mark.parametrize("leaf", [1, 2, 8])
def test_km_scale_grazing_rays_over_a_fine_plane_vs_brute_force(leaf):
    """Stress case for the culling rules: a sensor 2 m above a 2 km x 2 km ground tessellated into
    720 000 triangles, plus a near-coplanar second sheet 1 mm above part of it, and rays within 0.02 .. 3 degrees of the
    horizon -- Moeller-Trumbore's t is ill-conditioned there (absolute error of centimetres at several hundred
    metres).  The shipped traversal (tile lists forced on, every leaf size) must return exactly what the exhaustive
    variant returns, t and triangle id: a box or list entry is only ever culled on a distance bound that the exact
    test could not beat.  Reference behaviour being matched: rtcIntersect16's closest hit, SRC/EmbreeTracer.cpp:472-480."""
    flat = lambda X, Y: np.zeros_like(X)
    v0, t0 = synth.grid_mesh(600, 600, -1000.0, 1000.0, -1000.0, 1000.0, flat)
    v1, t1 = synth.grid_mesh(60, 60, 100.0, 700.0, -300.0, 300.0, lambda X, Y: np.full_like(X, 1e-3))
    elev = np.linspace(-0.02, -3.0, 24).astype(np.float32)
    res = []
    for mode in (6, 3):
        ctx = capi.Context(0, capi.FRAME_WORLD)
        ctx.set_option(capi.OPT_TRAVERSAL, mode)
        ctx.set_option(capi.OPT_LEAF_SIZE, leaf)
        ctx.set_option(capi.OPT_TILE_MIN_UNITS, 0)
        ctx.set_sensor(elev, 0.0, 360.0, 96, synth.yaw_quat(0.1), np.array([3.0, -7.0, 2.0], np.float32))
        for name, v, t in (("ground", v0, t0), ("sheet", v1, t1)):
            assert ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0]) >= 0
            assert ctx.update_geometry(name, v, t, IDENT) == 0
        assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0, ctx.last_error()
        assert ctx.trace_batch(0, 1, capi.TRACE_NO_POINTS) == 0, ctx.last_error()
        rc, t, tri = ctx.trace_raw(0)
        assert rc == 0
        res.append((t, tri))
        ctx.close()
    (t6, i6), (t3, i3) = res
    assert (i3 >= 0).sum() > 0.9 * i3.size and (t3[i3 >= 0] > 300.0).sum() > 100          # long grazing hits are in the set
    assert np.array_equal(i6, i3) and np.array_equal(t6.view(np.uint32), t3.view(np.uint32))


@pytest.mark.parametrize("watertight", [0, 1])
def test_second_return_vs_oracle(meshes, watertight):
    """LSB200_OPT_RETURNS = 2 (the any-hit multi-return the reference only plans, SRC/LidarShooter.hpp.in:87-93): per ray
    the next surface behind the first hit, bit-exact against the oracle -- XT-32 over ground + ben (rays that graze ben
    have the ground behind it) and a soup with duplicated triangles (second return at the SAME distance, larger id)."""
    args = xt32_sensor_args("0000")
    ctx = xt32_ctx(meshes, opts={capi.OPT_RETURNS: 2, capi.OPT_WATERTIGHT: watertight})
    assert ctx.commit_scene() == 0
    rc, t1, i1 = ctx.trace_raw()
    assert rc == 0
    rc, t2, i2 = ctx.fetch_second_returns()
    assert rc == 0, ctx.last_error()
    s = lso.Sensor(*args)
    with lso.watertight(bool(watertight)):
        verts, tris = lso.concat_scene(s, xt32_meshes(meshes), mode=0)
        bvh = lso.Bvh(verts, tris)
        dirs = lso.ray_dirs(s.tables)
        r1 = bvh.trace(np.zeros(3, np.float32), dirs)
        r2 = bvh.trace_second(np.zeros(3, np.float32), dirs, r1[0], r1[1])
        bvh.close()
    assert np.array_equal(i1, r1[1]) and np.array_equal(t1, r1[0])
    assert np.array_equal(i2, r2[1]) and np.array_equal(t2.view(np.uint32), r2[0].view(np.uint32))
    assert (i2 >= 0).sum() > 50 and np.all(t2[i2 >= 0] >= t1[i2 >= 0]) and np.all(i2[i1 < 0] == -1)
    ctx.close()

    rng = np.random.default_rng(3)
    n = 3000
    c = rng.uniform(-15, 15, (n, 1, 3)).astype(np.float32)
    v = (c + rng.normal(0, 1.2, (n, 3, 3)).astype(np.float32)).reshape(-1, 3)
    tr = np.arange(3 * n, dtype=np.uint32).reshape(-1, 3)
    tr = np.concatenate([tr, tr[:300]])
    elev = np.linspace(70, -70, 40).astype(np.float32)
    ctx = capi.Context(0)
    ctx.set_option(capi.OPT_RETURNS, 2)
    ctx.set_option(capi.OPT_WATERTIGHT, watertight)
    sargs = (elev, 0.0, 360.0, 120, (1, 0, 0, 0), (0, 0, 0))
    ctx.set_sensor(*sargs)
    assert ctx.add_geometry("soup", capi.GEOMETRY_TRIANGLE, v.shape[0], tr.shape[0]) >= 0
    assert ctx.update_geometry("soup", v, tr, IDENT) == 0
    assert ctx.commit_scene() == 0
    rc, t1, i1 = ctx.trace_raw()
    rc2, t2, i2 = ctx.fetch_second_returns()
    assert rc == 0 and rc2 == 0
    s = lso.Sensor(*sargs)
    with lso.watertight(bool(watertight)):
        verts, tris = lso.concat_scene(s, [(v, tr, IDENT)], mode=0)
        bvh = lso.Bvh(verts, tris)
        dirs = lso.ray_dirs(s.tables)
        r1 = bvh.trace(np.zeros(3, np.float32), dirs)
        r2 = bvh.trace_second(np.zeros(3, np.float32), dirs, r1[0], r1[1])
        bvh.close()
    assert np.array_equal(i1, r1[1]) and np.array_equal(i2, r2[1]) and np.array_equal(t2.view(np.uint32), r2[0].view(np.uint32))
    assert ((t2 == t1) & (i2 > i1) & (i1 >= 0)).sum() > 0
    ctx.close()
