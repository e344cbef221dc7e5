#!/usr/bin/env python
"""Stage times of the commit (full rebuild) on the bench scene: min over a few commits after a warm-up.
usage: build_stages.py [--reps N] [--small]   (LSB200_LIB selects an A/B build of the library)"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=6)
    ap.add_argument("--small", action="store_true")
    ap.add_argument("--fused", type=int, default=None, help="LSB200_OPT_FUSED_TREE (default: the library's)")
    args = ap.parse_args()
    from bench import build_workload, frame_affine
    from lidarshooter_b200 import capi
    meshes, poses, table = build_workload(args.small)
    ctx = capi.Context(0, capi.FRAME_WORLD)
    for gi, (name, v, t) in enumerate(meshes):
        assert ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0]) >= 0
        assert ctx.update_geometry(name, v, t, frame_affine(0, gi)) == 0
    ctx.set_option(capi.OPT_STAGE_TIMING, 1)
    if args.fused is not None:
        ctx.set_option(capi.OPT_FUSED_TREE, args.fused)
    keys = ("transform", "morton", "sort", "hierarchy", "fit")
    best = {k: 1e9 for k in keys}
    for r in range(args.reps + 2):
        assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0, ctx.last_error()
        st = ctx.stats()
        if r >= 2:
            for k in keys:
                best[k] = min(best[k], float(getattr(st, "ms_" + k)))
    best["total"] = sum(best[k] for k in keys)
    best["fused"] = args.fused
    best["lib"] = os.environ.get("LSB200_LIB", "")
    print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in best.items()}))
    ctx.close()


if __name__ == "__main__":
    main()
