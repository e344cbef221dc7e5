"""The CPU oracle against the reference's own pins (point counts) and its internal consistency.

Reference pins: 1668 (SRC/test/EmbreeTracer_test.cpp:134, SRC/test/OptixTracer_test.cpp:119),
1781 (SRC/test/OptixTracer_test.cpp:168), 0 for the empty scene (SRC/test/OptixTracer_test.cpp:307),
98 / 162 vertices / triangles for ground.stl (SRC/test/EmbreeTracer_test.cpp:89-90).
"""
import numpy as np
import pytest

from conftest import xt32_sensor_args
from oracle import lso


def _sensor(uid="0000"):
    return lso.Sensor(*xt32_sensor_args(uid))


def test_mesh_counts(meshes):
    assert meshes["ground_v"].shape == (98, 3) and meshes["ground_t"].shape == (162, 3)
    assert meshes["ben_v"].shape == (2823, 3) and meshes["ben_t"].shape == (5489, 3)


def test_ray_count_and_tables():
    s = _sensor()
    assert s.n_rays == 150 * 32                     # SRC/test/LidarDevice_test.cpp:58
    step = lso.azimuth_step(0.0, 360.0, 150)
    assert np.float32(0.0) + step * np.float32(149) == np.float32(360.0)
    d = lso.ray_dirs(s.tables)
    assert np.allclose(np.linalg.norm(d.astype(np.float64), axis=1), 1.0, atol=1e-6)
    # ray id = channel * H + azimuth; channel 15 is the horizontal one (elevation 0)
    assert abs(d[15 * 150, 2]) < 1e-7 and d[0, 2] > 0 and d[31 * 150, 2] < 0


@pytest.mark.parametrize("uid,n_ground,n_both", [("0000", 1668, 1781), ("0001", 1633, 1769)])
def test_reference_point_counts(meshes, uid, n_ground, n_both):
    s = _sensor(uid)
    g = (meshes["ground_v"], meshes["ground_t"], lso.IDENTITY_AFFINE)
    b = (meshes["ben_v"], meshes["ben_t"], lso.IDENTITY_AFFINE)
    r = lso.scan(s, [g])
    assert r["points"].shape[0] == n_ground
    r = lso.scan(s, [g, b])
    assert r["points"].shape[0] == n_both
    r = lso.scan(s, [])
    assert r["points"].shape[0] == 0 and (r["tri"] == -1).all()


def test_golden_vectors(meshes, golden):
    for uid in ("0000", "0001"):
        s = _sensor(uid)
        g = (meshes["ground_v"], meshes["ground_t"], lso.IDENTITY_AFFINE)
        b = (meshes["ben_v"], meshes["ben_t"], lso.IDENTITY_AFFINE)
        for tag, ms in (("ground", [g]), ("ground_ben", [g, b])):
            r = lso.scan(s, ms)
            assert np.array_equal(r["tri"], golden[f"l{uid}_{tag}_tri"])
            assert np.array_equal(r["t"], golden[f"l{uid}_{tag}_t"])


def test_bvh_equals_brute_on_random_soup():
    rng = np.random.default_rng(5)
    nT = 3000
    c = rng.uniform(-20, 20, (nT, 1, 3)).astype(np.float32)
    v = (c + rng.normal(0, 0.7, (nT, 3, 3)).astype(np.float32)).reshape(-1, 3)
    t = np.arange(3 * nT, dtype=np.uint32).reshape(-1, 3)
    d = rng.normal(size=(4000, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = rng.uniform(-5, 5, (4000, 3)).astype(np.float32)
    tb, ib = lso.trace_brute(v, t, o, d)
    bv = lso.Bvh(v, t)
    tv, iv = bv.trace(o, d)
    assert np.array_equal(ib, iv) and np.array_equal(tb, tv)
    assert (ib >= 0).sum() > 500


def test_world_frame_mode_agrees_with_sensor_frame(meshes):
    """Moving the rays instead of the scene changes rounding only: same hits away from edges, t within 1e-4 rel."""
    s = _sensor()
    g = (meshes["ground_v"], meshes["ground_t"], lso.IDENTITY_AFFINE)
    b = (meshes["ben_v"], meshes["ben_t"], lso.IDENTITY_AFFINE)
    r0, r1 = lso.scan(s, [g, b], mode=0), lso.scan(s, [g, b], mode=1)
    same = r0["tri"] == r1["tri"]
    assert (~same).sum() <= 8
    both = same & (r0["tri"] >= 0)
    assert np.allclose(r0["t"][both], r1["t"][both], rtol=1e-4, atol=0)


def test_point_record_layout(meshes):
    """SRC/XYZIRBytes.cpp:24-40: x y z | 4 zero | intensity 64.0 | ring i32 | 8 zero."""
    s = _sensor()
    r = lso.scan(s, [(meshes["ground_v"], meshes["ground_t"], lso.IDENTITY_AFFINE)])
    p = r["points"]
    assert p.shape[1] == 32
    f = p.view(np.float32).reshape(-1, 8)
    i = p.view(np.int32).reshape(-1, 8)
    assert (f[:, 4] == 64.0).all() and (i[:, 3] == 0).all() and (i[:, 6:] == 0).all()
    hit = np.flatnonzero(r["tri"] >= 0)
    assert np.array_equal(i[:, 5], (hit // s.H).astype(np.int32))
    d = lso.ray_dirs(s.tables)[hit]
    assert np.array_equal(f[:, :3], d * r["t"][hit, None])
    r0 = lso.scan(s, [(meshes["ground_v"], meshes["ground_t"], lso.IDENTITY_AFFINE)], ring_mode=1)
    assert (r0["points"].view(np.int32).reshape(-1, 8)[:, 5] == 0).all()    # Embree-path quirk, SRC/EmbreeTracer.cpp:333


def test_intersect_conventions():
    a, b, c = [0, 0, 1], [1, 0, 1], [0, 1, 1]
    hit, t = lso.intersect_one(a, b, c, [0.2, 0.2, 0], [0, 0, 1])
    assert hit and t == np.float32(1.0)
    hit, _ = lso.intersect_one(a, c, b, [0.2, 0.2, 0], [0, 0, 1])          # no back-face culling
    assert hit
    hit, _ = lso.intersect_one(a, b, c, [0.2, 0.2, 2], [0, 0, 1])          # behind the origin
    assert not hit
    hit, _ = lso.intersect_one(a, b, c, [0.2, 0.2, 1], [0, 0, 1])          # t = 0 is excluded (tnear strict)
    assert not hit
    hit, _ = lso.intersect_one(a, b, c, [0.2, 0.2, 0], [1, 0, 0])          # parallel: den = 0
    assert not hit
    hit, _ = lso.intersect_one(a, a, a, [0, 0, 0], [0, 0, 1])              # degenerate triangle
    assert not hit


def test_float64_second_opinion_xt32(meshes):
    """The float64 second opinion (textbook Moeller-Trumbore in double, oracle/grazing.py) over the reference's own
    scene: same 1668 / 1781 points as the fp32 oracle and the reference's tests, no grazing ray, range within 1e-6."""
    from oracle import grazing
    for uid, n_both in (("0000", 1781), ("0001", 1769)):
        s = lso.Sensor(*xt32_sensor_args(uid))
        r = lso.scan(s, [(meshes["ground_v"], meshes["ground_t"], lso.IDENTITY_AFFINE),
                         (meshes["ben_v"], meshes["ben_t"], lso.IDENTITY_AFFINE)])
        dirs = lso.ray_dirs(s.tables)
        t64, tri64 = lso.trace_brute_f64(r["verts"], r["tris"], np.zeros(3, np.float32), dirs)
        assert int((tri64 >= 0).sum()) == n_both
        rep = grazing.report(r["verts"], r["tris"], np.zeros(3, np.float32), dirs, r["t"], r["tri"], t64, tri64)
        assert rep["grazing_rays"] == 0 and rep["unexplained"] == 0, rep
        assert rep["max_rel_range_error_vs_f64"] < 1e-6


def test_grazing_report_finds_and_explains_edge_rays():
    """Rays aimed exactly at the shared diagonal / at the vertices of a tessellated quad, and rays skimming a plane
    edge-on: wherever fp32 and float64 disagree the report must name a cause; rays aimed at triangle interiors never
    disagree.  BVH and brute-force forms of the float64 tracer agree."""
    from oracle import grazing
    rng = np.random.default_rng(5)
    n = 24
    xs = np.linspace(-3.0, 3.0, n + 1, dtype=np.float32)
    X, Y = np.meshgrid(xs, xs, indexing="xy")
    Z = (0.3 * np.sin(1.7 * X) * np.cos(1.3 * Y)).astype(np.float32)
    verts = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(np.float32)
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="xy")
    v00 = (j * (n + 1) + i).astype(np.uint32)
    tris = np.concatenate([np.stack([v00, v00 + 1, v00 + n + 2], -1).reshape(-1, 3),
                           np.stack([v00, v00 + n + 2, v00 + n + 1], -1).reshape(-1, 3)]).astype(np.uint32)
    org = np.array([0.3, -0.2, 5.0], np.float32)
    # targets: vertices, edge midpoints (shared edges), random interior points
    a, b, c = verts[tris[:, 0]], verts[tris[:, 1]], verts[tris[:, 2]]
    w = rng.dirichlet((1, 1, 1), size=tris.shape[0]).astype(np.float32)
    interior = w[:, :1] * a + w[:, 1:2] * b + w[:, 2:] * c
    targets = np.concatenate([verts, 0.5 * (a + b), 0.5 * (b + c), 0.5 * (c + a), interior]).astype(np.float32)
    dirs = (targets - org).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True).astype(np.float32)
    t32, tri32 = lso.trace_brute(verts, tris, org, dirs)
    t64, tri64 = lso.trace_brute_f64(verts, tris, org, dirs)
    bvh = lso.Bvh(verts, tris)
    t64b, tri64b = bvh.trace_f64(org, dirs)
    bvh.close()
    assert np.array_equal(tri64, tri64b) and np.array_equal(t64, t64b)
    rep = grazing.report(verts, tris, org, dirs, t32, tri32, t64, tri64)
    assert rep["unexplained"] == 0, rep
    assert rep["grazing_rays"] > 0 and rep["grazing_rays"] == rep["edge"] + rep["tie"] + rep["edge_on"], rep
    n_int = interior.shape[0]
    assert np.array_equal(tri32[-n_int:], tri64[-n_int:])          # interior targets: never a disagreement
    assert rep["max_rel_range_error_vs_f64"] < 1e-5


def test_watertight_oracle_closes_leaks_on_edge_aimed_rays(meshes):
    """The optional watertight intersector of the oracle (Woop-Benthin-Wald 2013; include/lsb200.h: LSB200_OPT_WATERTIGHT):
    rays aimed at random points ON the shared edges and AT the interior vertices of a tessellated patch.  The reference's
    Moeller-Trumbore drops a share of them between two triangles (leaks); the watertight test drops none; the float64
    second opinion hits all of them.  On the reference's own scene both intersectors give the same 1781 points."""
    rng = np.random.default_rng(5)
    n = 48
    xs = np.linspace(-3.0, 3.0, n + 1, dtype=np.float32)
    X, Y = np.meshgrid(xs, xs, indexing="xy")
    Z = (0.3 * np.sin(1.7 * X) * np.cos(1.3 * Y)).astype(np.float32)
    verts = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(np.float32)
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="xy")
    v00 = (j * (n + 1) + i).astype(np.uint32)
    tris = np.concatenate([np.stack([v00, v00 + 1, v00 + n + 2], -1).reshape(-1, 3),
                           np.stack([v00, v00 + n + 2, v00 + n + 1], -1).reshape(-1, 3)]).astype(np.uint32)
    org = np.array([0.3, -0.2, 5.0], np.float32)
    a, b, c = verts[tris[:, 0]], verts[tris[:, 1]], verts[tris[:, 2]]
    lam = rng.uniform(0, 1, (tris.shape[0], 1)).astype(np.float32)
    pts = np.concatenate([verts, a + lam * (b - a), b + lam * (c - b), c + lam * (a - c)]).astype(np.float32)
    pts = pts[(np.abs(pts[:, 0]) < 2.9) & (np.abs(pts[:, 1]) < 2.9)]
    dirs = (pts - org).astype(np.float32)
    t_mt, tri_mt = lso.trace_brute(verts, tris, org, dirs)
    with lso.watertight():
        t_wt, tri_wt = lso.trace_brute(verts, tris, org, dirs)
        bvh = lso.Bvh(verts, tris)
        t_wb, tri_wb = bvh.trace(org, dirs)
        bvh.close()
    t64, tri64 = lso.trace_brute_f64(verts, tris, org, dirs)
    assert (tri64 >= 0).all()
    assert (tri_wt >= 0).all()                                   # watertight: no leak
    assert np.array_equal(tri_wt, tri_wb) and np.array_equal(t_wt, t_wb)
    leaks = int((tri_mt < 0).sum())
    assert leaks > 0                                             # the reference's intersector leaks on this ray set
    hit = tri_mt >= 0
    assert np.abs(t_wt[hit] - t_mt[hit]).max() < 1e-4
    assert np.abs(t_wt - t64).max() < 1e-4
    s = lso.Sensor(*xt32_sensor_args("0000"))
    ms = [(meshes["ground_v"], meshes["ground_t"], lso.IDENTITY_AFFINE), (meshes["ben_v"], meshes["ben_t"], lso.IDENTITY_AFFINE)]
    with lso.watertight():
        r = lso.scan(s, ms)
    assert r["points"].shape[0] == 1781
    assert lso.lib().lso_get_watertight() == 0


@pytest.mark.parametrize("n_rays", [1, 7, 8, 9, 4801])
def test_packet_tracer_is_bit_identical_to_the_scalar_walk(n_rays):
    """lso_trace_bvh_packet (8-ray AVX2 packets, the CPU baseline closer to rtcIntersect16 of
    SRC/EmbreeTracer.cpp:472-480) returns exactly the scalar oracle's (t, triangle id): ragged tails, misses,
    per-ray origins, ties between duplicated triangles."""
    rng = np.random.default_rng(n_rays)
    n_tris = 3000
    c = rng.uniform(-20, 20, (n_tris, 1, 3)).astype(np.float32)
    v = (c + rng.normal(0, 0.8, (n_tris, 3, 3)).astype(np.float32)).reshape(-1, 3)
    tr = np.arange(3 * n_tris, dtype=np.uint32).reshape(-1, 3)
    tr = np.concatenate([tr, tr[:200]])                       # duplicates: ties in t resolved by the smaller id
    bvh = lso.Bvh(v, tr)
    d = rng.normal(0, 1, (n_rays, 3)).astype(np.float32)
    d[::5, 2] = 0.0                                           # axis-parallel components: 0 * inf in the slab test
    for orgs in (np.zeros(3, np.float32), rng.uniform(-5, 5, (n_rays, 3)).astype(np.float32)):
        t0, i0 = bvh.trace(orgs, d)
        res = bvh.trace_packet(orgs, d)
        if res is None:
            pytest.skip("host without AVX2")
        assert np.array_equal(res[1], i0) and np.array_equal(res[0].view(np.uint32), t0.view(np.uint32))
    assert n_rays < 100 or (i0 >= 0).sum() > 100
    bvh.close()


def test_second_return_of_the_oracle_against_numpy_brute_force():
    """lso_trace_bvh_second: per ray the smallest (t, id) greater than the first hit -- checked against an exhaustive
    numpy evaluation built from lso.intersect_one-compatible brute force (trace_brute on the scene minus nothing:
    all hits enumerated per ray)."""
    rng = np.random.default_rng(5)
    n_tris = 400
    c = rng.uniform(-6, 6, (n_tris, 1, 3)).astype(np.float32)
    v = (c + rng.normal(0, 1.5, (n_tris, 3, 3)).astype(np.float32)).reshape(-1, 3)
    tr = np.arange(3 * n_tris, dtype=np.uint32).reshape(-1, 3)
    tr = np.concatenate([tr, tr[:40]])                        # duplicated triangles: second return at the SAME distance
    d = rng.normal(0, 1, (300, 3)).astype(np.float32)
    o = np.zeros(3, np.float32)
    bvh = lso.Bvh(v, tr)
    t1, i1 = bvh.trace(o, d)
    t2, i2 = bvh.trace_second(o, d, t1, i1)
    # exhaustive: every (ray, triangle) hit through the single-triangle entry
    for r in range(d.shape[0]):
        hits = []
        for g in range(tr.shape[0]):
            hit, t = lso.intersect_one(v[tr[g, 0]], v[tr[g, 1]], v[tr[g, 2]], o, d[r])
            if hit:
                hits.append((np.float32(t), g))
        hits.sort()
        assert (hits[0] if hits else (np.float32(np.inf), -1)) == (t1[r], i1[r]) or (not hits and i1[r] < 0)
        want = hits[1] if len(hits) > 1 else (np.float32(np.inf), -1)
        assert (t2[r], i2[r]) == want, r
    assert (i2 >= 0).sum() > 20 and (t2[:] == t1[:]).sum() > 0
    bvh.close()
