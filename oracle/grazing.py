"""Grazing-ray report: fp32 oracle vs the float64 second opinion.  TEST INFRASTRUCTURE ONLY (see lsb_oracle.c).

north_star: "hit/miss mask and hit triangle index bit-exact except where a ray grazes a shared edge (counted and
reported)".  The reference's intersector is Embree's single-precision Moeller-Trumbore (SRC/EmbreeTracer.cpp:472-480);
its hit/miss decision on a ray that passes within rounding distance of a triangle edge depends on the exact operation
order and instruction selection (FMA, rcp) of the Embree build, which cannot be restated here.  What CAN be measured
is the set of rays on which ANY correct single-precision intersector may legitimately differ from exact arithmetic:

    grazing rays = rays whose (hit mask, triangle id) from the fp32 oracle differ from textbook Moeller-Trumbore
                   evaluated in float64 on the same fp32 vertices and rays (lso_trace_bvh_f64).

Every such ray is then classified, in float64, against the triangles the two answers name:

    edge      the ray passes within k fp32-ulp (conditioned, see below) of an edge of one of the two triangles
    tie       both triangles contain the hit point and their distances agree to 1e-6 relative (shared edge / vertex,
              duplicate or coplanar overlapping triangles): the two answers differ only in the tie-break
    edge_on   the ray is (nearly) parallel to the plane of one of the triangles
    unexplained   none of the above -- must be 0

Conditioning: the fp32 test computes U = dot(cross(C, D), e2) with C = v0 - o; its rounding error is a few
2^-24 |C| |D| |e|, so in barycentric units (U / den) the edge distance that single precision can resolve is
ulp_b = 2^-24 * M * |e|max * |d| / |det| with M = max(|C|, |v0|, |o|) and det = e1 . (d x e2).  The report gives, per
grazing ray, the distance to the nearest edge in units of ulp_b and asserts it stays below `k_ulp`.
"""
import numpy as np

from . import lso


def _tri_geometry(verts, tris, org, dirs, ids):
    v = np.asarray(verts, np.float64)
    t = np.asarray(tris, np.int64)[ids]
    a, b, c = v[t[:, 0]], v[t[:, 1]], v[t[:, 2]]
    e1, e2 = b - a, c - a
    o = np.asarray(org, np.float64).reshape(-1, 3)
    C = a - o
    d = np.asarray(dirs, np.float64)
    det = np.einsum("ij,ij->i", e1, np.cross(d, e2))
    emax = np.maximum(np.linalg.norm(e1, axis=1), np.maximum(np.linalg.norm(e2, axis=1), np.linalg.norm(c - b, axis=1)))
    M = np.maximum(np.linalg.norm(C, axis=1), np.maximum(np.linalg.norm(a, axis=1), np.linalg.norm(o, axis=1)))
    dn = np.linalg.norm(d, axis=1)
    with np.errstate(divide="ignore", invalid="ignore"):
        ulp_b = 2.0 ** -24 * M * emax * dn / np.abs(det)
        sin_plane = np.abs(det) / (np.linalg.norm(e1, axis=1) * np.linalg.norm(e2, axis=1) * dn)
    return ulp_b, sin_plane


def _edge_distance_ulps(verts, tris, org, dirs, ids):
    """Per ray: |signed distance to the triangle's boundary| in barycentric units / ulp_b, the f64 t, sin(ray, plane).
    Rows with id < 0 give inf."""
    n = len(ids)
    out_u = np.full(n, np.inf)
    out_t = np.full(n, np.nan)
    out_s = np.full(n, np.inf)
    ok = ids >= 0
    if ok.any():
        sel = np.nonzero(ok)[0]
        uvtd = lso.bary_f64(verts, tris, org, dirs[sel], ids[sel])
        u, v, t = uvtd[:, 0], uvtd[:, 1], uvtd[:, 2]
        # signed distance to the boundary: > 0 outside.  Inside: the nearest edge is min(u, v, 1 - u - v) away.
        outside = np.maximum(np.maximum(-u, -v), u + v - 1.0)
        ulp_b, sin_plane = _tri_geometry(verts, tris, org, dirs[sel], ids[sel])
        with np.errstate(divide="ignore", invalid="ignore"):
            out_u[sel] = np.abs(outside) / ulp_b
        out_t[sel] = t
        out_s[sel] = sin_plane
    return out_u, out_t, out_s


def report(verts, tris, org, dirs, t32, tri32, t64, tri64, k_ulp=64.0):
    """Compares the fp32 answer (t32, tri32) with the float64 second opinion (t64, tri64) over the same rays.
    Returns a dict; `unexplained` must be 0."""
    dirs = np.asarray(dirs, np.float32).reshape(-1, 3)
    tri32 = np.asarray(tri32, np.int32)
    tri64 = np.asarray(tri64, np.int32)
    differ = tri32 != tri64                       # covers the mask too (-1 = miss)
    idx = np.nonzero(differ)[0]
    res = {"rays": int(dirs.shape[0]), "hits_fp32": int((tri32 >= 0).sum()), "hits_f64": int((tri64 >= 0).sum()),
           "grazing_rays": int(idx.size), "mask_flips": int(((tri32 >= 0) != (tri64 >= 0)).sum()),
           "edge": 0, "tie": 0, "edge_on": 0, "unexplained": 0, "max_edge_ulps": 0.0, "k_ulp": float(k_ulp)}
    both = (tri32 >= 0) & (tri64 >= 0) & ~differ
    if both.any():
        rel = np.abs(t32[both].astype(np.float64) - t64[both]) / np.maximum(t64[both], 1e-30)
        res["max_rel_range_error_vs_f64"] = float(rel.max())
    if idx.size == 0:
        return res
    d = dirs[idx]
    u32, tt32, s32 = _edge_distance_ulps(verts, tris, org, d, tri32[idx])
    u64, tt64, s64 = _edge_distance_ulps(verts, tris, org, d, tri64[idx])
    near = np.minimum(u32, u64)
    edge = near <= k_ulp
    both_hit = (tri32[idx] >= 0) & (tri64[idx] >= 0)
    with np.errstate(invalid="ignore"):
        tie = both_hit & (np.abs(tt32 - tt64) <= 1e-6 * np.maximum(np.abs(tt64), 1e-30))
    edge_on = np.minimum(s32, s64) < 1e-5
    cls_edge = edge
    cls_tie = ~cls_edge & tie
    cls_on = ~cls_edge & ~cls_tie & edge_on
    unexpl = ~(cls_edge | cls_tie | cls_on)
    res.update(edge=int(cls_edge.sum()), tie=int(cls_tie.sum()), edge_on=int(cls_on.sum()), unexplained=int(unexpl.sum()),
               max_edge_ulps=float(near[cls_edge].max()) if cls_edge.any() else 0.0)
    if unexpl.any():
        res["unexplained_rays"] = [int(x) for x in idx[unexpl][:16]]
        res["unexplained_edge_ulps"] = [float(x) for x in near[unexpl][:16]]
    return res


def second_opinion(verts, tris, org, dirs, t32, tri32, bvh=None, k_ulp=64.0):
    """Runs the float64 tracer over the rays and returns report()."""
    own = bvh is None
    if own:
        bvh = lso.Bvh(verts, tris)
    t64, tri64 = bvh.trace_f64(org, dirs)
    if own:
        bvh.close()
    return report(verts, tris, org, dirs, np.asarray(t32, np.float32), tri32, t64, tri64, k_ulp)
