#!/usr/bin/env python
"""Minimal driver for ncu: builds the bench scene once and launches the traversal kernel a few times.
usage: profile_trace.py [--mode M] [--leaf L] [--top T] [--launches N] [--small]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", type=int, default=None)
    ap.add_argument("--leaf", type=int, default=None)
    ap.add_argument("--top", type=int, default=None)
    ap.add_argument("--launches", type=int, default=3)
    ap.add_argument("--small", action="store_true")
    ap.add_argument("--flags", type=int, default=0)
    ap.add_argument("--max-range", type=float, default=0.0, help="detection radius for every sensor (0 = none)")
    ap.add_argument("--tile-entries", type=int, default=None)
    ap.add_argument("--tile-visits", type=int, default=None)
    args = ap.parse_args()
    from bench import build_workload, frame_affine
    from lidarshooter_b200 import capi
    meshes, poses, table = build_workload(args.small)
    elev, b, e, H = table
    ctx = capi.Context(0, capi.FRAME_WORLD)
    if args.mode is not None:
        ctx.set_option(capi.OPT_TRAVERSAL, args.mode)
    if args.leaf is not None:
        ctx.set_option(capi.OPT_LEAF_SIZE, args.leaf)
    if args.top is not None:
        ctx.set_option(capi.OPT_TOP_NODES, args.top)
    for k, (q, t) in enumerate(poses):
        ctx.set_sensor(elev, b, e, H, q, t, index=k)
    if args.max_range > 0:
        for k in range(len(poses)):
            assert ctx.set_sensor_limits(args.max_range, index=k) == 0
    for gi, (name, v, t) in enumerate(meshes):
        assert ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0]) >= 0
        assert ctx.update_geometry(name, v, t, frame_affine(0, gi)) == 0
    assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0, ctx.last_error()
    for _ in range(args.launches):
        if args.tile_entries is not None:
            ctx.set_option(capi.OPT_TILE_ENTRIES, args.tile_entries)      # also drops the cached lists
        if args.tile_visits is not None:
            ctx.set_option(capi.OPT_TILE_VISITS, args.tile_visits)
        assert ctx.trace_batch(0, len(poses), args.flags) == 0, ctx.last_error()
    ctx.synchronize()
    print("ok")
    ctx.close()


if __name__ == "__main__":
    main()
