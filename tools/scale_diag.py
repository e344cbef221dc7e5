#!/usr/bin/env python
"""Per-rank view of the resident bench step under torchrun (why does max-over-ranks ms_commit grow at 8 GPUs?).
Every rank writes gpurun_out/scale_diag_rank<r>.json: per-step commit / trace milliseconds (CUDA events), host microseconds
per call, CPU affinity, SM clock samples.  Evidence tool, not the product path."""
import json
import os
import subprocess
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import build_workload, frame_affine  # noqa: E402
from lidarshooter_b200 import capi  # noqa: E402

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
meshes, poses, table = build_workload(False)
elev, b, e, H = table
stream = torch.cuda.Stream()
ctx = capi.Context(local, capi.FRAME_WORLD, stream=stream.cuda_stream)
for k, (q, t) in enumerate(poses):
    ctx.set_sensor(elev, b, e, H, q, t, index=k)
names = []
keep = []
for gi, (name, v, t) in enumerate(meshes):
    pv = torch.from_numpy(v).pin_memory(); keep.append(pv)
    ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0])
    ctx.update_geometry(name, pv.numpy(), t, frame_affine(0, gi))
    names.append(name)


def step(frame):
    for gi, name in enumerate(names):
        ctx.update_transform(name, frame_affine(frame, gi))
    assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0


def sync_all():
    stream.synchronize(); torch.cuda.synchronize()
    if world > 1:
        dist.barrier(); torch.cuda.synchronize()


for w in range(3):
    step(w * world + rank); ctx.trace_batch(0, len(poses), 0)
sync_all()
ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
host = []
smi = subprocess.Popen(["nvidia-smi", "-i", str(local), "--query-gpu=clocks.sm,clocks.mem,power.draw,clocks_throttle_reasons.active",
                        "--format=csv,noheader", "-lms", "20"], stdout=subprocess.PIPE, text=True)
for k in range(steps):
    t0 = time.perf_counter()
    ev[k][0].record(stream)
    step((3 + k) * world + rank)
    ev[k][1].record(stream)
    t1 = time.perf_counter()
    ctx.trace_batch(0, len(poses), 0)
    ev[k][2].record(stream)
    t2 = time.perf_counter()
    host.append(((t1 - t0) * 1e6, (t2 - t1) * 1e6))
sync_all()
smi.terminate()
clk = smi.stdout.read().strip().splitlines()
out = {"rank": rank, "world": world, "affinity": sorted(os.sched_getaffinity(0)),
       "ms_commit": [round(ev[k][0].elapsed_time(ev[k][1]), 4) for k in range(steps)],
       "ms_trace": [round(ev[k][1].elapsed_time(ev[k][2]), 4) for k in range(steps)],
       "ms_total": round(ev[0][0].elapsed_time(ev[-1][2]), 4),
       "host_us_commit": [round(h[0], 1) for h in host], "host_us_trace": [round(h[1], 1) for h in host],
       "smi": clk[:3] + clk[-3:], "smi_samples": len(clk)}
ctx.set_option(capi.OPT_STAGE_TIMING, 1)
step(99 * world + rank); ctx.trace_batch(0, len(poses), 0)
st = ctx.stats()
out["stages"] = {k: round(float(getattr(st, "ms_" + k)), 4) for k in ("transform", "morton", "sort", "hierarchy", "fit", "trace")}
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"scale_diag_rank{rank}.json"), "w"))
print(json.dumps({k: out[k] for k in ("rank", "ms_total", "stages")}))
if world > 1:
    dist.destroy_process_group()
