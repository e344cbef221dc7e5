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
    ap.add_argument("--product", action="store_true", help="load liblsb200.so (the shipped build) instead of the lab build")
    ap.add_argument("--rebuild-each", action="store_true", help="full commit (rebuild) before every trace launch, like a bench step")
    ap.add_argument("--profiler-range", action="store_true", help="cudaProfilerStart/Stop around the launches after the first "
                    "(use with ncu --profile-from-start off)")
    args = ap.parse_args()
    from bench import build_workload, frame_affine
    from lidarshooter_b200 import capi
    meshes, poses, table = build_workload(args.small)
    elev, b, e, H = table
    ctx = capi.Context(0, capi.FRAME_WORLD, lab=not args.product)
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
    rt = None
    if args.profiler_range:
        import ctypes
        rt = ctypes.CDLL("libcudart.so")
    for it in range(args.launches):
        if rt is not None and it == 1:
            ctx.synchronize(); rt.cudaProfilerStart()
        if args.rebuild_each:
            for gi, (name, v, t) in enumerate(meshes):
                ctx.update_transform(name, frame_affine(it + 1, gi))
            assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0, ctx.last_error()
        if args.tile_entries is not None:
            ctx.set_option(capi.OPT_TILE_ENTRIES, args.tile_entries)      # also drops the cached lists
        if args.tile_visits is not None:
            ctx.set_option(capi.OPT_TILE_VISITS, args.tile_visits)
        assert ctx.trace_batch(0, len(poses), args.flags) == 0, ctx.last_error()
    ctx.synchronize()
    if rt is not None:
        rt.cudaProfilerStop()
    print("ok")
    ctx.close()


if __name__ == "__main__":
    main()
