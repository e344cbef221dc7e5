#!/usr/bin/env python
"""Host-side cost of one bench step (the calls are asynchronous: their duration is what the host thread pays to enqueue a frame).
Prints microseconds per call, median over the steps, with the GPU drained before every step so that no call blocks on a full queue."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import build_workload, frame_affine  # noqa: E402
from lidarshooter_b200 import capi  # noqa: E402

meshes, poses, table = build_workload(False)
elev, b, e, H = table
ctx = capi.Context(0, capi.FRAME_WORLD)
for k, (q, t) in enumerate(poses):
    ctx.set_sensor(elev, b, e, H, q, t, index=k)
names = []
for gi, (name, v, t) in enumerate(meshes):
    ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0])
    ctx.update_geometry(name, v, t, frame_affine(0, gi))
    names.append(name)
T = {"affines (numpy)": [], "update_transform x9": [], "commit_scene": [], "trace_batch": []}
for f in range(12):
    t0 = time.perf_counter()
    A = [frame_affine(f, gi) for gi in range(len(names))]
    t1 = time.perf_counter()
    for gi, name in enumerate(names):
        ctx.update_transform(name, A[gi])
    t2 = time.perf_counter()
    ctx.commit_scene(capi.COMMIT_REBUILD)
    t3 = time.perf_counter()
    ctx.trace_batch(0, len(poses), 0)
    t4 = time.perf_counter()
    ctx.synchronize()
    if f >= 2:
        for k, d in zip(T, (t1 - t0, t2 - t1, t3 - t2, t4 - t3)):
            T[k].append(d * 1e6)
for k, v in T.items():
    print(f"{k:22s} {np.median(v):8.1f} us (min {min(v):.1f}, max {max(v):.1f})")
