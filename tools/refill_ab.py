#!/usr/bin/env python
"""A/B of the lane-refill mask (LSB200_OPT_LANE_REFILL) on the bench workload (configs[3]): per mask, the trace stage of
a bench step (rebuild, tile lists, k_trace) timed with the library's CUDA events, min / mean over --reps steps, and a
CRC of two sensors' per-ray (t, triangle id) compared with the first mask's.  LSB200_LIB selects an A/B build.
usage: refill_ab.py [--masks 0,7,15] [--reps 6] [--small]"""
import argparse
import json
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--masks", default="0,1,2,4,8,7,15")
    ap.add_argument("--reps", type=int, default=6)
    ap.add_argument("--small", action="store_true")
    args = ap.parse_args()
    from bench import build_workload, frame_affine
    from lidarshooter_b200 import capi
    meshes, poses, table = build_workload(args.small)
    elev, b, e, H = table
    ns = len(poses)
    ctx = capi.Context(0, capi.FRAME_WORLD, lab=True)      # run_phase_refill lives in the lab library
    if args.small:
        ctx.set_option(capi.OPT_TILE_MIN_UNITS, 0)
    for k, (q, t) in enumerate(poses):
        ctx.set_sensor(elev, b, e, H, q, t, index=k)
    for gi, (name, v, t) in enumerate(meshes):
        assert ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0]) >= 0
        assert ctx.update_geometry(name, v, t, frame_affine(0, gi)) == 0
    ctx.set_option(capi.OPT_STAGE_TIMING, 1)
    base = None
    for mask in [int(x) for x in args.masks.split(",")]:
        ctx.set_option(capi.OPT_LANE_REFILL, mask)
        tr, tl = [], []
        for r in range(args.reps + 2):
            for gi, (name, v_, t_) in enumerate(meshes):
                ctx.update_transform(name, frame_affine(r + 3, gi))
            assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0, ctx.last_error()
            assert ctx.trace_batch(0, ns, capi.TRACE_XYZI16) == 0, ctx.last_error()
            st = ctx.stats()
            if r >= 2:
                tr.append(float(st.ms_trace)); tl.append(float(st.ms_tile_lists))
        sig = []
        for s in (0, ns - 1):
            rc, t, tri = ctx.trace_raw(s)
            assert rc == 0
            sig.append((zlib.crc32(t.tobytes()), zlib.crc32(tri.tobytes()), int((tri >= 0).sum())))
        if base is None:
            base = sig
        k = [a - b for a, b in zip(tr, tl)]
        print(json.dumps({"refill_mask": mask, "ms_stage_min": round(min(tr), 4), "ms_stage_mean": round(float(np.mean(tr)), 4),
                          "ms_k_trace_min": round(min(k), 4), "ms_k_trace_mean": round(float(np.mean(k)), 4),
                          "same_results_as_first": sig == base, "lib": os.environ.get("LSB200_LIB", "")}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
