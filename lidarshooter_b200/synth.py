"""Synthetic workloads of BASELINE.json (configs 3-5): procedurally tessellated scenes and LiDAR tables.

Pure numpy, seeded, deterministic; shared by tests/ and bench.py.  Nothing here is on the product's compute path:
it only manufactures the INPUTS (vertex / index arrays, elevation tables, sensor poses) that are handed to the
C ABI exactly like the reference hands pcl::PolygonMesh blobs to ITracer::updateGeometry.
Definitions follow SURVEY.md section 8(d); the PCG32 streams named there are realised with numpy's PCG64
`default_rng(seed)` (documented in DESIGN.md).
"""
import numpy as np


def grid_mesh(nx, ny, x0, x1, y0, y1, zfun):
    """(nx x ny) cells, 2 triangles per cell -> ((nx+1)(ny+1) verts, 2 nx ny tris)."""
    xs = np.linspace(x0, x1, nx + 1, dtype=np.float64)
    ys = np.linspace(y0, y1, ny + 1, dtype=np.float64)
    X, Y = np.meshgrid(xs, ys, indexing="xy")            # (ny+1, nx+1)
    Z = zfun(X, Y)
    verts = np.stack([X, Y, Z], axis=-1).reshape(-1, 3).astype(np.float32)
    i = np.arange(nx, dtype=np.uint32)[None, :]
    j = np.arange(ny, dtype=np.uint32)[:, None]
    v00 = (j * (nx + 1) + i).astype(np.uint32)
    v10, v01, v11 = v00 + 1, v00 + (nx + 1), v00 + (nx + 2)
    t0 = np.stack([v00, v10, v11], axis=-1)
    t1 = np.stack([v00, v11, v01], axis=-1)
    tris = np.stack([t0, t1], axis=2).reshape(-1, 3).astype(np.uint32)
    return verts, tris


def heightfield(nx=1000, ny=500, half_x=250.0, half_y=125.0, seed=42):
    """Config 3 scene: 1000 x 500 cells over [-250,250] x [-125,125] m = 1 000 000 triangles, 501 501 vertices;
    z = sum_k a_k sin(f_k x + p_k) sin(g_k y + q_k), 8 terms."""
    rng = np.random.default_rng(seed)
    a = rng.uniform(0.1, 1.5, 8)
    f = rng.uniform(0.01, 0.2, 8)
    g = rng.uniform(0.01, 0.2, 8)
    p = rng.uniform(0, 2 * np.pi, 8)
    q = rng.uniform(0, 2 * np.pi, 8)

    def z(X, Y):
        out = np.zeros_like(X)
        for k in range(8):
            out += a[k] * np.sin(f[k] * X + p[k]) * np.sin(g[k] * Y + q[k])
        return out

    return grid_mesh(nx, ny, -half_x, half_x, -half_y, half_y, z)


def _box(cx, cy, wx, wy, h, n=10):
    """Axis-aligned box without a bottom: 5 faces, each n x n cells x 2 = 200 triangles -> 1000 triangles."""
    vs, ts, off = [], [], 0
    u = np.linspace(0.0, 1.0, n + 1)
    U, V = np.meshgrid(u, u, indexing="xy")
    i = np.arange(n, dtype=np.uint32)[None, :]
    j = np.arange(n, dtype=np.uint32)[:, None]
    v00 = (j * (n + 1) + i).astype(np.uint32)
    cell = np.stack([np.stack([v00, v00 + 1, v00 + n + 2], -1), np.stack([v00, v00 + n + 2, v00 + n + 1], -1)], 2).reshape(-1, 3)
    x0, x1, y0, y1 = cx - wx / 2, cx + wx / 2, cy - wy / 2, cy + wy / 2
    faces = [
        (x0 + U * wx, y0 + V * wy, np.full_like(U, h)),           # roof
        (x0 + U * wx, np.full_like(U, y0), V * h),                # -y wall
        (x0 + U * wx, np.full_like(U, y1), V * h),                # +y wall
        (np.full_like(U, x0), y0 + U * wy, V * h),                # -x wall
        (np.full_like(U, x1), y0 + U * wy, V * h),                # +x wall
    ]
    for X, Y, Z in faces:
        vs.append(np.stack([X, Y, Z], -1).reshape(-1, 3))
        ts.append(cell + np.uint32(off))
        off += (n + 1) * (n + 1)
    return np.concatenate(vs).astype(np.float32), np.concatenate(ts).astype(np.uint32)


CITY_PITCH = 2000.0 / 90.0      # street-grid pitch of city(): 90 x 90 lots over 2 km


def city(n_boxes=8000, ground_cells=1000, extent=2000.0, seed=2024, n_groups=8, lots=90):
    """Config 4 scene: ground grid 1000 x 1000 cells over 2 km x 2 km (2 000 000 triangles) + 8000 box buildings x
    1000 triangles = 10 000 000 triangles.  Buildings stand on a lots x lots street grid (one per lot, the lots left
    empty chosen by the seeded stream), centre jittered by +-2 m, footprint 8-16 m, height 5-80 m, so streets stay
    open and no two buildings overlap (built-up fraction ~30 %).  Returns a list of (name, verts, tris): 'ground' +
    n_groups building meshes (buildings dealt round-robin), i.e. 9 geometries."""
    rng = np.random.default_rng(seed)
    half = extent / 2
    pitch = extent / lots
    ground = grid_mesh(ground_cells, ground_cells, -half, half, -half, half, lambda X, Y: np.zeros_like(X))
    n_boxes = min(n_boxes, lots * lots)
    lot = rng.permutation(lots * lots)[:n_boxes]
    li, lj = np.divmod(lot, lots)
    cx = (lj + 0.5) * pitch - half + rng.uniform(-2, 2, n_boxes)
    cy = (li + 0.5) * pitch - half + rng.uniform(-2, 2, n_boxes)
    fmax = min(16.0, pitch - 6.0)
    wx = rng.uniform(fmax / 2, fmax, n_boxes)
    wy = rng.uniform(fmax / 2, fmax, n_boxes)
    hh = rng.uniform(5, 80, n_boxes)
    # one template box, scaled/translated per instance (fast, vectorised)
    tv, tt = _box(0.0, 0.0, 1.0, 1.0, 1.0)
    out = [("ground", ground[0], ground[1])]
    n_groups = max(1, min(n_groups, n_boxes)) if n_boxes else 0
    for gi in range(n_groups):
        ids = np.arange(gi, n_boxes, n_groups)
        m = ids.size
        V = tv[None, :, :] * np.stack([wx[ids], wy[ids], hh[ids]], -1)[:, None, :].astype(np.float32)
        V[:, :, 0] += cx[ids, None].astype(np.float32)
        V[:, :, 1] += cy[ids, None].astype(np.float32)
        T = tt[None, :, :] + (np.arange(m, dtype=np.uint32) * np.uint32(tv.shape[0]))[:, None, None]
        out.append((f"buildings_{gi}", V.reshape(-1, 3).astype(np.float32), T.reshape(-1, 3).astype(np.uint32)))
    return out


def lidar_table(channels=128, az_count=3601, elev_hi=15.0, elev_lo=-25.0):
    """Config 3 device: 128 channels, elevations linspace(+15, -25), azimuth 0..360 in 3601 steps (0.1 deg)."""
    elev = np.linspace(elev_hi, elev_lo, channels).astype(np.float32)
    return elev, 0.0, 360.0, az_count


XT32_ELEV = np.arange(15, -17, -1, dtype=np.float32)      # config/hesai-pandar-XT-32-lidar_0000.json:77-81


def yaw_quat(yaw):
    """(w, x, y, z) of a rotation about +z."""
    return np.array([np.cos(yaw / 2), 0.0, 0.0, np.sin(yaw / 2)], np.float32)


def sensor_lattice(n_side=8, spacing=10 * CITY_PITCH, z=5.0, seed=7):
    """Config 4 devices: n_side^2 sensors on a lattice centred on the origin, yaw from the seeded stream.  With the
    default spacing (10 lots) every sensor stands on a street crossing of city(): lot corners are at integer
    multiples of the pitch and (k - 3.5) * 10 is an integer."""
    rng = np.random.default_rng(seed)
    yaws = rng.uniform(0, 2 * np.pi, n_side * n_side)
    c = (np.arange(n_side) - (n_side - 1) / 2) * spacing
    poses = []
    for k in range(n_side * n_side):
        i, j = divmod(k, n_side)
        # a small offset so that sensors do not sit exactly above grid lines of the ground mesh
        poses.append((yaw_quat(yaws[k]), np.array([c[j] + 0.137, c[i] + 0.211, z], np.float32)))
    return poses


def trajectory_poses(segments, period=0.1):
    """The headless stand-in for joystick -> AffineMesh integration (SRC/AffineMesh.cpp:108-128), one pose per
    tick: lin += Rz(ang.z) Ry(ang.y) Rx(ang.x) (linear * dt); ang += angular * dt.  segments: [(dt, linear[3],
    angular[3])].  Returns [(translation[3], euler_xyz[3])] in float32."""
    lin = np.zeros(3, np.float32)
    ang = np.zeros(3, np.float32)
    out = []
    for seg_dt, linear, angular in segments:
        n = int(round(seg_dt / period))
        linear = np.asarray(linear, np.float32)
        angular = np.asarray(angular, np.float32)
        for _ in range(n):
            cx, sx = np.cos(ang[0]), np.sin(ang[0])
            cy, sy = np.cos(ang[1]), np.sin(ang[1])
            cz, sz = np.cos(ang[2]), np.sin(ang[2])
            Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]], np.float32)
            Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]], np.float32)
            Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]], np.float32)
            lin = (lin + (Rz @ Ry @ Rx) @ (linear * np.float32(period))).astype(np.float32)
            ang = (ang + angular * np.float32(period)).astype(np.float32)
            out.append((lin.copy(), ang.copy()))
    return out
