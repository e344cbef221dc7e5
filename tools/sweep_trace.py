#!/usr/bin/env python
"""Traversal design sweep on the bench workload (configs[3], 10 M-triangle city, 64 devices): for every
(Morton scaling, leaf size, top-block size, kernel variant) the trace launch is timed with CUDA events and an
instrumented launch reports nodes / leaves / triangles per ray.  Every variant's per-ray (t, triangle id) of two
sensors is compared bit-for-bit with the first variant's.  Output: one JSON object per line (gpurun_out/...).

Evidence tool, not the product path: it only calls the C ABI."""
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
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep_trace.jsonl"))
    ap.add_argument("--small", action="store_true")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--leaves", default="1,2,4,8")
    ap.add_argument("--modes", default="0,4,1,2")
    ap.add_argument("--tops", default="1024")
    ap.add_argument("--morton", default="0,1")
    ap.add_argument("--chain", default="1")
    ap.add_argument("--tiles", default="8")
    ap.add_argument("--layouts", default="0")
    ap.add_argument("--tile-entries", default="16", help="per-tile entry list sizes (0 = off)")
    ap.add_argument("--tile-visits", default="0", help="node visits per tile list (0 = default)")
    ap.add_argument("--super", default="1", help="two-level lists (supertiles) on / off")
    ap.add_argument("--aspects", default="50")
    ap.add_argument("--scene", default="city", choices=["city", "heightfield"])
    ap.add_argument("--sway", action="store_true", help="with --recommit: move the geometries as bench.py's frames do")
    ap.add_argument("--recommit", action="store_true", help="rebuild the scene before every timed launch, as a bench step does (cold L2)")
    args = ap.parse_args()
    from bench import build_workload, frame_affine
    from lidarshooter_b200 import capi

    meshes, poses, table = build_workload(args.small)
    if args.scene == "heightfield":      # BASELINE configs[2]: 1M-triangle heightfield, 4 sensors above it
        from lidarshooter_b200 import synth
        v, t = synth.heightfield()
        meshes = [("terrain", v, t)]
        poses = [(synth.yaw_quat(0.3 * k), np.array([60.0 * k - 90.0, 20.0 * k - 30.0, 4.0], np.float32)) for k in range(4)]
    elev, b, e, H = table
    R = elev.size * H
    ns = len(poses)
    ctx = capi.Context(0, capi.FRAME_WORLD, lab=True)
    for k, (q, t) in enumerate(poses):
        ctx.set_sensor(elev, b, e, H, q, t, index=k)
    for gi, (name, v, t) in enumerate(meshes):
        assert ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0]) >= 0
        assert ctx.update_geometry(name, v, t, frame_affine(0, gi)) == 0
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    fout = open(args.out, "w")
    base_sig = None
    check_sensors = (0, ns - 1)
    ctx.set_option(capi.OPT_STAGE_TIMING, 1)
    for mu, layout, aspect in [(int(x), int(y), int(z)) for x in args.morton.split(",") for y in args.layouts.split(",") for z in args.aspects.split(",")]:
        for leaf in [int(x) for x in args.leaves.split(",")]:
            for top in [int(x) for x in args.tops.split(",")]:
                ctx.set_option(capi.OPT_MORTON_UNIFORM, mu)
                ctx.set_option(capi.OPT_MORTON_LAYOUT, layout)
                ctx.set_option(capi.OPT_MORTON_ASPECT, aspect)
                ctx.set_option(capi.OPT_LEAF_SIZE, leaf)
                ctx.set_option(capi.OPT_TOP_NODES, top)
                assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0, ctx.last_error()
                st = ctx.stats()
                build_ms = {k: round(float(getattr(st, "ms_" + k)), 4) for k in
                            ("transform", "morton", "sort", "hierarchy", "fit", "toplevel")}
                for mode, chain, tile, te, tv in [(int(x), int(y), int(z), int(u), int(w)) for x in args.modes.split(",") for y in args.chain.split(",")
                                                  for z in args.tiles.split(",") for u in args.tile_entries.split(",") for w in args.tile_visits.split(",")]:
                  for sup in [int(x) for x in args.super.split(",")]:
                    ctx.set_option(capi.OPT_TRAVERSAL, mode)
                    ctx.set_option(capi.OPT_TILE_VISITS, tv)
                    ctx.set_option(capi.OPT_SUPER_TILES, sup)
                    ctx.set_option(capi.OPT_ORIGIN_CHAIN, chain)
                    ctx.set_option(capi.OPT_TILE_WIDTH, tile)
                    ctx.set_option(capi.OPT_TRACE_STATS, 0)
                    ms = []
                    for r in range(args.reps + 1):
                        ctx.set_option(capi.OPT_TILE_ENTRIES, te)      # also drops the cached lists: every timed launch rebuilds them
                        if args.recommit:
                            if args.sway:
                                for gi, (name, v_, t_) in enumerate(meshes):
                                    ctx.update_transform(name, frame_affine(r + 3, gi))
                            assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0, ctx.last_error()
                        assert ctx.trace_batch(0, ns, 0) == 0, ctx.last_error()
                        ms.append(float(ctx.stats().ms_trace))
                    ms = ms[1:]
                    ctx.set_option(capi.OPT_TRACE_STATS, 1)
                    assert ctx.trace_batch(0, ns, capi.TRACE_NO_POINTS) == 0, ctx.last_error()
                    cnt = ctx.trace_counters()
                    ctx.set_option(capi.OPT_TRACE_STATS, 0)
                    sig = []
                    for s in check_sensors:
                        rc, t, tri = ctx.trace_raw(s)
                        assert rc == 0
                        sig.append((zlib.crc32(t.tobytes()), zlib.crc32(tri.tobytes()), int((tri >= 0).sum())))
                    if base_sig is None:
                        base_sig = sig
                    rays = max(cnt["rays"], 1)
                    row = {"morton_uniform": mu, "morton_layout": layout, "morton_aspect": aspect, "leaf": leaf, "top": top, "mode": mode, "chain": chain, "tile_w": tile, "tile_entries": te, "tile_visits": tv, "super_tiles": sup, "super_visits_per_tile": cnt["super_visits"] / max(cnt["units_uniform"], 1),
                           "tile_visits_per_tile": cnt["tile_visits"] / max(cnt["units_uniform"], 1), "tile_entries_per_tile": cnt["tile_entries"] / max(cnt["units_uniform"], 1),
                           "tile_fallbacks": cnt["tile_fallbacks"], "tile_empty": cnt["tile_empty"], "chain_tests_per_ray": cnt["chain_tests"] / rays,
                           "ms_trace_min": min(ms), "ms_trace_mean": float(np.mean(ms)),
                           "mrays_per_s": ns * R / (min(ms) * 1e-3) / 1e6,
                           "nodes_per_ray": cnt["nodes"] / rays, "leaves_per_ray": cnt["leaves"] / rays,
                           "tris_per_ray": cnt["tris"] / rays, "max_stack": cnt["max_stack"], "culled_pops_per_ray": cnt["culled_pops"] / rays,
                           "units_uniform": cnt["units_uniform"], "units_mixed": cnt["units_mixed"],
                           "hit_fraction": cnt["hits"] / rays, "same_results_as_first": sig == base_sig,
                           "grid": int(ctx.stats().trace_grid), "build_ms": build_ms}
                    print(json.dumps(row), flush=True)
                    fout.write(json.dumps(row) + "\n")
                    fout.flush()
    ctx.close()


if __name__ == "__main__":
    main()
