#!/usr/bin/env python
"""ONE process driving N GPUs through the C ABI's multi-GPU entry (lsb200_clone_scene + lsb200_run_batch) on the bench
workload (BASELINE configs[3]) -- the cross-check VERDICT r01 asked for: the whole-job rate through one process must
equal the torchrun (one process per GPU) figure of bench.py.

  frames   LSB200_SHARD_FRAMES : GPU g takes frames g, g + N, ... (throughput view; every GPU rebuilds its own frames)
  sensors  LSB200_SHARD_SENSORS: every GPU takes every frame and 64 / N of the sensors (latency view: every GPU pays
           the full rebuild, so a frame finishes in ~ commit + trace / N -- the Amdahl term of the replicated build)

Prints one JSON line.  usage: multi_gpu_batch.py [--gpus N] [--frames F] [--record xyzi16|range|none] [--small]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=0)
    ap.add_argument("--frames", type=int, default=16, help="frames per GPU in the frame-sharded run")
    ap.add_argument("--small", action="store_true")
    args = ap.parse_args()
    import torch
    from bench import build_workload, frame_affine
    from lidarshooter_b200 import capi
    n = args.gpus or torch.cuda.device_count()
    meshes, poses, table = build_workload(args.small)
    elev, b, e, H = table
    R, S = elev.size * H, len(poses)
    ctx0 = capi.Context(0, capi.FRAME_WORLD)
    for k, (q, t) in enumerate(poses):
        ctx0.set_sensor(elev, b, e, H, q, t, index=k)
    for gi, (name, v, t) in enumerate(meshes):
        assert ctx0.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0]) >= 0
        assert ctx0.update_geometry(name, v, t, frame_affine(0, gi)) == 0
    t0 = time.perf_counter()
    ctxs = [ctx0] + [ctx0.clone(g) for g in range(1, n)]
    t_clone = time.perf_counter() - t0
    out = {"tool": "multi_gpu_batch (one process, lsb200_run_batch)", "gpus": n, "clone_scene_s": t_clone,
           "n_triangles": int(sum(m[2].shape[0] for m in meshes)), "devices": S, "rays_per_device": R}
    G = len(meshes)
    for label, flags, rec in (("xyzi16", capi.TRACE_XYZI16, 16), ("range_image", capi.TRACE_RANGE_IMAGE, 4)):
        # pinned output: two frames' worth per GPU, NUMA-placed by the library; frames reuse them round-robin
        bufs = [[c.alloc_host(S * R * rec) for _ in range(2)] for c in ctxs]
        for sharding, name in ((capi.SHARD_FRAMES, "frames"), (capi.SHARD_SENSORS, "sensors")):
            F = args.frames * n if sharding == capi.SHARD_FRAMES else args.frames
            A = np.stack([np.stack([frame_affine(f, gi) for gi in range(G)]) for f in range(F)])
            ptrs = []
            for f in range(F):
                if sharding == capi.SHARD_FRAMES:
                    g, slot = f % n, (f // n) % 2
                    base = int(bufs[g][slot].ctypes.data)
                    ptrs.append([base + k * R * rec for k in range(S)])
                else:
                    row = []
                    for k in range(S):
                        g = min(n - 1, next(gg for gg in range(n) if k < S * (gg + 1) // n))
                        row.append(int(bufs[g][f % 2].ctypes.data) + k * R * rec)
                    ptrs.append(row)
            for rep in range(2):                       # first pass = warm-up
                t0 = time.perf_counter()
                rc, cnt = capi.run_batch(ctxs, A, ptrs, [R] * S, first=0, count=S, commit_mode=capi.COMMIT_REBUILD, flags=flags,
                                         sharding=sharding)
                dt = time.perf_counter() - t0
                assert rc == 0, [c.last_error() for c in ctxs]
            out[f"{label}_{name}"] = {"frames": F, "ms_per_frame": dt / F * 1e3, "mrays_per_s": F * S * R / dt / 1e6,
                                      "points_per_frame": int(cnt.sum()) // F}
        for c, bb in zip(ctxs, bufs):
            for x in bb:
                c.free_host(x)
    for c in ctxs[1:]:
        c.close()
    ctx0.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
