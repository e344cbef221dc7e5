#!/usr/bin/env python
"""bench.py -- the trace-path benchmark of lidarshooter-b200 (contract: see DESIGN.md "Measurement").

Workload (BASELINE.json configs[3], the configuration the metric is quoted on): the synthetic 10 M-triangle city,
64 LiDAR devices (128 channels x 3601 azimuth steps = 460 928 rays each, 29.5 M rays per frame), world-frame mode.
One STEP = one frame on every rank: per-geometry affine update -> vertex transform -> full LBVH rebuild (Morton,
radix sort, Karras, AABB fit, top-level block) -> closest-hit trace of all 64 devices in one launch -> stream
compaction into 32-byte PointCloud2 records.  Ranks shard the trajectory FRAMES (rank r works on frame
step * N + r); the scene is replicated per GPU and there is no data-path collective ("scaling": "weak").

  value      Mrays/s, whole job, inputs resident in HBM, device-timed (CUDA events on the launching stream)
  e2e        the same through the C ABI with HOST buffers: H2D of every mesh's vertices from pinned memory and D2H
             of every device's point cloud inside the timed region
  roofline   the traversal kernel against the measured HBM peak, algorithmic bytes per SURVEY.md section 8(d)
  cpu_baseline / --impl reference   the CPU oracle (oracle/, a port: Embree is not installable here) on the host cores

Prints ONE JSON line on rank 0.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_SENSORS = 64
METRIC = "Mrays/s (10M-tri city, 64 devices x 460928 rays, per-frame rebuild + trace + compaction)"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def b_ray(n_tris, hit_fraction):
    """Algorithmic bytes per ray: one root-to-leaf path of 64-B nodes + one 48-B triangle + output."""
    return 64.0 * math.ceil(math.log2(max(n_tris, 2))) + 48.0 + 32.0 * hit_fraction


def build_workload(small=False):
    from lidarshooter_b200 import synth
    if small:
        meshes = synth.city(n_boxes=380, ground_cells=200, extent=20 * synth.CITY_PITCH, n_groups=4, lots=20)
        poses = synth.sensor_lattice(2, 10 * synth.CITY_PITCH)
        table = synth.lidar_table(32, 361)
    else:
        meshes = synth.city()
        poses = synth.sensor_lattice(8)
        table = synth.lidar_table()
    return meshes, poses, table


def frame_affine(frame, gi):
    """Per-frame rigid motion of geometry gi (buildings sway by centimetres so that every frame is new work)."""
    A = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)
    if gi > 0:
        A[3] = np.float32(0.05 * math.sin(0.37 * frame + gi))
        A[7] = np.float32(0.05 * math.cos(0.23 * frame + 2 * gi))
    return A


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu_index):
        self.rows, self.proc, self.thread = [], None, None
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def wait_first(self, timeout_s=5.0):
        """Blocks until nvidia-smi has delivered its first sample: its start-up (NVML init over every GPU of the box, hundreds
        of milliseconds on an 8-GPU node, under driver locks that stall kernel launches of every rank) must not overlap the
        timed region -- at 8 ranks it cost 0.17 ms per step of a 27 ms region (profiles/scale_diag_r02_8gpu.txt)."""
        t0 = time.time()
        while self.proc and not self.rows and time.time() - t0 < timeout_s:
            time.sleep(0.01)

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ts, line in self.rows:
            if ts < t0 - 0.05 or ts > t1 + 0.15:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except Exception:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:      # region shorter than one sample: take whatever was seen
            for ts, line in self.rows[-3:]:
                f = [x.strip() for x in line.split(",")]
                try:
                    sm.append(float(f[0])); mx.append(float(f[1]))
                except Exception:
                    pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_trace_sample(meshes, poses, table, n_threads=None, sensors=(0,), repeats=1, affines=None, gpu_results=None):
    """The CPU oracle on the host cores: world-frame transform (per-geometry `affines`, identity when None) +
    binned-SAH BVH build once, then closest-hit trace of the rays of `sensors`.
    gpu_results: {sensor: (t, tri)} from lsb200_trace_raw of the SAME frame -- compared ray by ray (the parity check of
    the benchmarked configuration), and the float64 second opinion is run over the same rays (grazing report).
    Returns dict(mrays_trace, t_build, t_trace, rays, cores[, parity])."""
    from oracle import lso, grazing
    lso.build()
    if n_threads:
        lso.set_threads(n_threads)
    cores = lso.max_threads()
    elev, b, e, H = table
    t0 = time.perf_counter()
    ident = lso.IDENTITY_AFFINE
    if affines is None:
        affines = [ident] * len(meshes)
    s0 = lso.Sensor(elev, b, e, H, poses[0][0], poses[0][1])
    verts, tris = lso.concat_scene(s0, [(v, t, a) for (_, v, t), a in zip(meshes, affines)], mode=1)
    t1 = time.perf_counter()
    bvh = lso.Bvh(verts, tris)
    t2 = time.perf_counter()
    rays, t_trace, hits = 0, 0.0, 0
    t_packet, packet_rays, packet_same = 0.0, 0, True
    parity = None
    for rep in range(repeats):
        for si in sensors:
            s = lso.Sensor(elev, b, e, H, poses[si][0], poses[si][1])
            dirs = lso.rotate_dirs(s.R, lso.ray_dirs(s.tables))
            ta = time.perf_counter()
            t, tri = bvh.trace(s.T, dirs)
            t_trace += time.perf_counter() - ta
            tp = time.perf_counter()
            pk = bvh.trace_packet(s.T, dirs)              # 8-ray AVX2 packets: the baseline closer to rtcIntersect16
            if pk is not None:
                t_packet += time.perf_counter() - tp
                packet_rays += dirs.shape[0]
                packet_same = packet_same and np.array_equal(pk[1], tri) and np.array_equal(pk[0].view(np.uint32), t.view(np.uint32))
            rays += dirs.shape[0]
            hits += int((tri >= 0).sum())
            if rep == 0 and gpu_results is not None and si in gpu_results:
                gt, gtri = gpu_results[si]
                g = grazing.second_opinion(verts, tris, s.T, dirs, gt, gtri, bvh=bvh)
                if parity is None:
                    parity = {"against": "oracle CPU BVH (bit-exact) + float64 second opinion (oracle/grazing.py)",
                              "sensors": [], "rays": 0, "mismatch_mask": 0, "mismatch_tri": 0, "mismatch_t": 0,
                              "grazing_rays": 0, "grazing_unexplained": 0, "grazing_by_cause": {"edge": 0, "tie": 0, "edge_on": 0},
                              "max_edge_ulps": 0.0, "max_rel_range_error_vs_f64": 0.0}
                parity["sensors"].append(int(si))
                parity["rays"] += int(dirs.shape[0])
                parity["mismatch_mask"] += int(((gtri >= 0) != (tri >= 0)).sum())
                parity["mismatch_tri"] += int((gtri != tri).sum())
                parity["mismatch_t"] += int((gt.view(np.uint32) != t.view(np.uint32)).sum())
                parity["grazing_rays"] += g["grazing_rays"]
                parity["grazing_unexplained"] += g["unexplained"]
                for k in ("edge", "tie", "edge_on"):
                    parity["grazing_by_cause"][k] += g[k]
                parity["max_edge_ulps"] = max(parity["max_edge_ulps"], g["max_edge_ulps"])
                parity["max_rel_range_error_vs_f64"] = max(parity["max_rel_range_error_vs_f64"],
                                                           g.get("max_rel_range_error_vs_f64", 0.0))
    bvh.close()
    return {"mrays_trace": rays / t_trace / 1e6, "t_transform": t1 - t0, "t_build": t2 - t1, "t_trace": t_trace,
            "rays": rays, "hits": hits, "cores": cores, "n_tris": int(tris.shape[0]), "parity": parity,
            "mrays_packet": (packet_rays / t_packet / 1e6) if packet_rays else None, "t_packet": t_packet,
            "packet_same_as_scalar": bool(packet_same) if packet_rays else None}


def xt32_scan_leg(local, reps=1000, cpu=True):
    """BASELINE configs[0]: Hesai XT-32 lidar_0000, one scan of ground + ben (5 651 triangles, 4 800
    rays), sensor-frame mode like the reference.  Tiny and latency-bound, so it is reported as us/scan and scans/s:
      e2e      = per scan what MeshProjector does per frame (SRC/MeshProjector.cpp:446-471): updateGeometry of both
                 meshes from host vertices, commitScene (full rebuild, as rtcCommitScene does), traceScene into a host
                 cloud of 32-byte records -- wall clock of the blocking C-ABI calls;
      resident = transform update + rebuild + trace with no host copies, CUDA events;
      cpu      = the oracle's scan (same arithmetic as the reference's CPU path) on the host cores."""
    import torch
    from lidarshooter_b200 import capi
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import GOLDEN, xt32_sensor_args
    z = np.load(os.path.join(GOLDEN, "meshes.npz"))
    args = xt32_sensor_args("0000")
    stream = torch.cuda.Stream()
    ctx = capi.Context(local, capi.FRAME_SENSOR, stream=stream.cuda_stream)
    ctx.set_sensor(*args)
    ident = np.asarray([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)
    host = {}
    for name in ("ground", "ben"):
        v = torch.from_numpy(np.ascontiguousarray(z[name + "_v"], np.float32)).pin_memory()
        t = np.ascontiguousarray(z[name + "_t"], np.uint32)
        host[name] = (v, t)
        assert ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0]) >= 0
        assert ctx.update_geometry(name, v.numpy(), t, ident) == 0
    R = ctx.ray_count(0)
    out = torch.empty((R, 32), dtype=torch.uint8).pin_memory().numpy()

    def scan_e2e():
        for name, (v, t) in host.items():
            ctx.update_geometry(name, v.numpy(), None, ident)
        rc = ctx.commit_scene(capi.COMMIT_REBUILD)
        rc2, pts = ctx.trace_scene(0, 0, out=out)
        return rc, rc2, pts

    for _ in range(20):
        rc, rc2, pts = scan_e2e()
    assert rc == 0 and rc2 == 0 and pts.shape[0] == 1781, (rc, rc2, pts.shape)    # SRC/test/OptixTracer_test.cpp:168
    t0 = time.perf_counter()
    for _ in range(reps):
        scan_e2e()
    us_e2e = (time.perf_counter() - t0) / reps * 1e6
    launches0 = int(ctx.stats().kernel_launches)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stream.synchronize()
    e0.record(stream)
    for _ in range(reps):
        ctx.update_transform("ben", ident)
        ctx.commit_scene(capi.COMMIT_REBUILD)
        ctx.trace_device(0, 0)
    e1.record(stream)
    stream.synchronize()
    us_res = e0.elapsed_time(e1) * 1e3 / reps
    launches = (int(ctx.stats().kernel_launches) - launches0) / reps
    ctx.close()

    # throughput with one tracer per device (the reference runs one tracer per LiDAR, SRC/mainwindow.cpp:256-274):
    # K independent contexts on K streams, scans issued round-robin from this one host thread, nothing read back
    K = 16
    streams = [torch.cuda.Stream() for _ in range(K)]
    ctxs = []
    for k in range(K):
        cx = capi.Context(local, capi.FRAME_SENSOR, stream=streams[k].cuda_stream)
        cx.set_sensor(*args)
        for name, (v, t) in host.items():
            cx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0])
            cx.update_geometry(name, v.numpy(), t, ident)
        ctxs.append(cx)

    def round_robin(n):
        for _ in range(n):
            for cx in ctxs:
                cx.update_transform("ben", ident)
                cx.commit_scene(capi.COMMIT_REBUILD)
                cx.trace_device(0, 0)

    round_robin(5)
    torch.cuda.synchronize()
    rr = max(reps // K, 10)
    t0 = time.perf_counter()
    round_robin(rr)
    torch.cuda.synchronize()
    us_conc = (time.perf_counter() - t0) / (rr * K) * 1e6
    n_check = ctxs[-1].fetch_points(0)[1].shape[0]
    assert n_check == 1781, n_check
    for cx in ctxs:
        cx.close()
    leg = {"workload": "configs[0]: XT-32 lidar_0000, ground + ben, 5651 triangles, 4800 rays, sensor frame, rebuild per scan",
           "points": 1781, "reps": reps,
           "e2e_us_per_scan": us_e2e, "e2e_scans_per_s": 1e6 / us_e2e, "e2e_mrays_per_s": R / us_e2e,
           "resident_us_per_scan": us_res, "resident_scans_per_s": 1e6 / us_res, "launches_per_scan": launches,
           "concurrent_tracers": K, "concurrent_us_per_scan": us_conc, "concurrent_scans_per_s": 1e6 / us_conc}
    if cpu:
        from oracle import lso
        lso.build()
        sens = lso.Sensor(*args)
        ms = [(z[n + "_v"], z[n + "_t"], lso.IDENTITY_AFFINE) for n in ("ground", "ben")]
        best = None
        for use_bvh in (False, True):
            lso.scan(sens, ms, use_bvh=use_bvh)
            t0 = time.perf_counter()
            n = 5
            for _ in range(n):
                lso.scan(sens, ms, use_bvh=use_bvh)
            us = (time.perf_counter() - t0) / n * 1e6
            if best is None or us < best[0]:
                best = (us, "bvh" if use_bvh else "brute")
        leg["cpu_us_per_scan"] = best[0]
        leg["cpu_scans_per_s"] = 1e6 / best[0]
        leg["cpu_kind"] = f"oracle port ({best[1]}), {lso.max_threads()} host threads, transform + build + trace + packing"
    return leg


def other_configs_leg(local, rank, world, n_scans=60):
    """BASELINE configs 2, 3 and 5 as extra keys of the bench line -- the headline stays
    configs[3].  Each is measured end to end through the public call a user of that configuration makes:
      c2  two XT-32 devices (lidar_0000 + lidar_0001), `ben` along config/trajectory.json over the static ground, through the
          drop-in B200Tracer adapter (updateGeometry x 2, commitScene = refit, traceScene into the PointCloud2); scans/s
      c3  one 128 x 3601 device over the 1 M-triangle heightfield: adapter call sequence (refit per scan) and the C ABI with
          a full rebuild per scan
      c5  1 M-triangle heightfield + 16 `ben` on looped trajectories, FULL rebuild every frame, frames sharded over the
          ranks (frame-major): frames/s of the whole job through lsb200_run_frames with host clouds (16-byte records)
    c2 / c3 run on rank 0 only (single-device configurations); c5 on every rank."""
    import torch
    import torch.distributed as dist
    from lidarshooter_b200 import capi, hostapi, synth
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import GOLDEN
    out = {}
    z = np.load(os.path.join(GOLDEN, "meshes.npz"))
    ident = np.asarray([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)
    if rank == 0:
        # ---- c2 ----------------------------------------------------------------------------------------------------
        poses = hostapi.trajectory_poses(os.path.join(GOLDEN, "trajectory.json"), 0.1)        # 150 ticks
        ground = hostapi.Mesh(os.path.join(GOLDEN, "ground.stl"))
        ben = hostapi.Mesh(arrays=(z["ben_v"], z["ben_t"]))
        tracers = []
        for uid in ("0000", "0001"):
            dev = hostapi.LidarDevice(os.path.join(GOLDEN, f"xt32_lidar_{uid}.json"))
            tr = hostapi.B200Tracer(dev, local)
            tr.addGeometry("ground", 0, 98, 162); tr.addGeometry("ben", 0, 2823, 5489)
            tracers.append((dev, tr))

        def c2_pass():
            pts = 0
            for k in range(poses.shape[0]):
                for _, tr in tracers:
                    tr.updateGeometry("ground", ground, translation=(0, 0, 0), rotation=(0, 0, 0))
                    tr.updateGeometry("ben", ben, translation=poses[k, :3], rotation=poses[k, 3:])
                    tr.commitScene()
                    tr.traceScene(k)
                    pts += tr.getTraceCloudWidth()
            return pts
        c2_pass()
        t0 = time.perf_counter()
        pts = c2_pass()
        dt = time.perf_counter() - t0
        scans = 2 * poses.shape[0]
        out["c2"] = {"workload": "configs[1]: lidar_0000 + lidar_0001, 150 trajectory frames, per-frame transforms + BVH refit, "
                                 "through lidarshooter::B200Tracer (sensor frame, host PointCloud2)",
                     "scans": scans, "us_per_scan": dt / scans * 1e6, "scans_per_s": scans / dt, "mrays_per_s": scans * 4800 / dt / 1e6,
                     "points_total": int(pts), "reference_design_rate_scans_per_s": 20.0}
        for dev, tr in tracers:
            tr.close(); dev.close()
        # ---- c3 ----------------------------------------------------------------------------------------------------
        hv, ht = synth.heightfield()
        elev, b, e, H = synth.lidar_table()
        R = elev.size * H
        dev = hostapi.LidarDevice(synthetic=(elev, b, e, H, (1, 0, 0, 0), (0.0, 0.0, 2.0)))
        tr = hostapi.B200Tracer(dev, local)
        terrain = hostapi.Mesh(arrays=(hv, ht))
        tr.addGeometry("terrain", 0, hv.shape[0], ht.shape[0])
        n3 = 30

        def c3_adapter(n):
            for k in range(n):
                tr.updateGeometry("terrain", terrain, translation=(0.01 * k, 0, 0), rotation=(0, 0, 0))
                tr.commitScene()
                tr.traceScene(k)
        c3_adapter(3)
        t0 = time.perf_counter()
        c3_adapter(n3)
        dt_ad = (time.perf_counter() - t0) / n3
        hits3 = tr.getTraceCloudWidth()
        tr.close(); dev.close()
        stream = torch.cuda.Stream()
        ctx = capi.Context(local, capi.FRAME_SENSOR, stream=stream.cuda_stream)
        ctx.set_sensor(elev, b, e, H, (1, 0, 0, 0), (0.0, 0.0, 2.0))
        ctx.add_geometry("terrain", capi.GEOMETRY_TRIANGLE, hv.shape[0], ht.shape[0])
        ctx.update_geometry("terrain", hv, ht, ident)
        for mode_name, mode in (("rebuild", capi.COMMIT_REBUILD), ("refit", capi.COMMIT_REFIT)):
            for k in range(3):
                ctx.commit_scene(capi.COMMIT_REBUILD if k == 0 else mode); ctx.trace_device(0, 0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            stream.synchronize()
            e0.record(stream)
            for k in range(n3):
                A = ident.copy(); A[3] = 0.01 * k
                ctx.update_transform("terrain", A)
                ctx.commit_scene(mode)
                ctx.trace_device(0, 0)
            e1.record(stream)
            stream.synchronize()
            out.setdefault("c3", {})[f"resident_{mode_name}_ms_per_scan"] = e0.elapsed_time(e1) / n3
        ctx.close()
        out["c3"].update({"workload": "configs[2]: 128 ch x 3601 az (460 928 rays), 1 M-triangle heightfield, sensor frame",
                          "adapter_ms_per_scan": dt_ad * 1e3, "adapter_mrays_per_s": R / dt_ad / 1e6, "adapter_scans_per_s": 1.0 / dt_ad,
                          "adapter_call": "B200Tracer::updateGeometry (1 M-triangle mesh from host, pageable) + commitScene (refit) + "
                                          "traceScene (32-byte records into the std::vector of the PointCloud2)",
                          "resident_rebuild_mrays_per_s": R / (out["c3"]["resident_rebuild_ms_per_scan"] * 1e-3) / 1e6,
                          "points": int(hits3)})
    # ---- c5 (every rank) ------------------------------------------------------------------------------------------------
    hv, ht = synth.heightfield()
    bv, bt = z["ben_v"], z["ben_t"]
    loop = synth.trajectory_poses([(10.0, (0, 5, 0), (0, 0, 1)), (5.0, (1, 0, 0), (0, 0, -1))])
    n_inst = 16
    elev, b, e, H = synth.lidar_table()
    R = elev.size * H
    ctx = capi.Context(local, capi.FRAME_WORLD)
    ctx.set_sensor(elev, b, e, H, (1, 0, 0, 0), (0.0, 0.0, 4.0))
    ctx.add_geometry("terrain", capi.GEOMETRY_TRIANGLE, hv.shape[0], ht.shape[0]); ctx.update_geometry("terrain", hv, ht, ident)
    base = [np.array([(i % 4 - 1.5) * 12.0, (i // 4 - 1.5) * 12.0, 3.0], np.float32) for i in range(n_inst)]
    for i in range(n_inst):
        ctx.add_geometry(f"ben{i:02d}", capi.GEOMETRY_TRIANGLE, bv.shape[0], bt.shape[0]); ctx.update_geometry(f"ben{i:02d}", bv, bt, ident)
    frames = list(range(rank, n_scans * world, world))               # frame-major sharding of the 4096-frame batch's head
    A = np.zeros((len(frames), 1 + n_inst, 12), np.float32)
    for k, f in enumerate(frames):
        A[k, 0] = ident
        for i in range(n_inst):
            lin, ang = loop[(f + 9 * i) % len(loop)]
            A[k, 1 + i] = capi.affine_from_components(lin * np.float32(0.2) + base[i], ang)
    bufs = [ctx.alloc_host(R * 16) for _ in range(2)]
    ptrs = [[int(bufs[k % 2].ctypes.data)] for k in range(len(frames))]
    ctx.run_frames(A[:2], ptrs[:2], [R], first=0, count=1, commit_mode=capi.COMMIT_REBUILD, flags=capi.TRACE_XYZI16)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    rc, cnt = ctx.run_frames(A, ptrs, [R], first=0, count=1, commit_mode=capi.COMMIT_REBUILD, flags=capi.TRACE_XYZI16)
    dt = time.perf_counter() - t0
    assert rc == 0, ctx.last_error()
    if world > 1:
        tt = torch.tensor([dt], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
    n_tris5 = int(ctx.stats().n_triangles)
    for bb in bufs:
        ctx.free_host(bb)
    ctx.close()
    if rank == 0:
        total = len(frames) * world
        out["c5"] = {"workload": "configs[4]: 1 M-triangle heightfield + 16 ben on looped trajectories, full LBVH rebuild every frame, "
                                 "128 x 3601 rays, frames sharded over the ranks; head of the 4096-frame batch",
                     "frames": total, "n_triangles": n_tris5, "frames_per_s": total / dt, "ms_per_frame_per_gpu": dt / len(frames) * 1e3,
                     "mrays_per_s": total * R / dt / 1e6, "seconds_for_4096_frames": 4096 / (total / dt),
                     "api": "lsb200_run_frames, 16-byte records to pinned host memory"}
    return out


_RESULT_FD = None


def emit_result(line):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def run_reference(args):
    """--impl reference: the reference's CPU algorithm for the path.  Embree 3.13.4 (the reference's intersector) is
    a closed binary that is neither in /root/reference nor installable here, so this is the oracle PORT (oracle/),
    all host threads.  Each step traces one device's 460 928 rays against the 10 M-triangle CPU BVH."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import lso
    lso.build()
    # all the host threads the process may use -- torchrun exports OMP_NUM_THREADS=1 to its workers
    try:
        lso.set_threads(len(os.sched_getaffinity(0)))
    except AttributeError:
        lso.set_threads(os.cpu_count() or 1)
    meshes, poses, table = build_workload(args.small)
    elev, b, e, H = table
    ident = lso.IDENTITY_AFFINE
    s0 = lso.Sensor(elev, b, e, H, poses[0][0], poses[0][1])
    verts, tris = lso.concat_scene(s0, [(v, t, ident) for _, v, t in meshes], mode=1)
    t0 = time.perf_counter()
    bvh = lso.Bvh(verts, tris)
    t_build = time.perf_counter() - t0
    cores = lso.max_threads()
    R = elev.size * H
    times = []
    for step in range(args.warmup + args.steps):
        si = step % len(poses)
        s = lso.Sensor(elev, b, e, H, poses[si][0], poses[si][1])
        ta = time.perf_counter()
        dirs_s = lso.ray_dirs(s.tables)
        dirs = lso.rotate_dirs(s.R, dirs_s)
        pk = bvh.trace_packet(s.T, dirs)                  # 8-ray AVX2 packets (bit-identical to the scalar walk)
        t, tri = pk if pk is not None else bvh.trace(s.T, dirs)
        packets = pk is not None
        pts = lso.pack_points(dirs_s, t, tri, H)
        dt = time.perf_counter() - ta
        if step >= args.warmup:
            times.append(dt)
    total = sum(times)
    val = R * len(times) / total / 1e6
    sample = (f"each step = ray generation + closest-hit trace ({'8-ray AVX2 packets' if packets else 'scalar rays: host without AVX2'}, "
              f"all host threads) + point packing of ONE of the {len(poses)} devices "
              f"({R} rays) against the {tris.shape[0]}-triangle scene; CPU BVH (binned SAH) built once outside the "
              f"timed region in {t_build:.1f} s (the reference rebuilds per device per frame: not charged)")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args.small), "n_triangles": int(tris.shape[0]), "rays_per_step": R},
            "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit_result(line)


def workload_name(small):
    return ("SMALL smoke city (not the benchmark)" if small else
            "configs[3]: synthetic 10M-triangle city, 64 LiDAR devices (128ch x 3601 az), world frame")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--small", action="store_true", help="tiny scene for smoke runs (NOT the benchmark)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle / float64 comparison of the benchmarked frame")
    ap.add_argument("--no-xt32", action="store_true", help="skip the XT-32 scan leg (profiling runs of the main step)")
    ap.add_argument("--no-other-configs", action="store_true", help="skip the configs[1] / [2] / [4] legs")
    ap.add_argument("--traversal", type=int, default=None)
    ap.add_argument("--top-nodes", type=int, default=None)
    ap.add_argument("--leaf", type=int, default=None)
    ap.add_argument("--commit", default="rebuild", choices=["rebuild", "auto"])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    # stdout carries exactly ONE line, the JSON result: everything libraries print to file descriptor 1 meanwhile
    # (NCCL's version banner, for one) is sent to stderr, and the line is written to the saved descriptor at the end
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from lidarshooter_b200 import capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL's banner ("NCCL version ...") goes to stdout by default: keep stdout for the ONE JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    meshes, poses, table = build_workload(args.small)
    elev, b, e, H = table
    R = elev.size * H
    n_sens = len(poses)
    stream = torch.cuda.Stream()
    ctx = capi.Context(local, capi.FRAME_WORLD, stream=stream.cuda_stream)
    if args.traversal is not None:
        ctx.set_option(capi.OPT_TRAVERSAL, args.traversal)
    if args.top_nodes is not None:
        ctx.set_option(capi.OPT_TOP_NODES, args.top_nodes)
    if args.leaf is not None:
        ctx.set_option(capi.OPT_LEAF_SIZE, args.leaf)
    for k, (q, t) in enumerate(poses):
        ctx.set_sensor(elev, b, e, H, q, t, index=k)
    # pinned host copies of the meshes (what a host application owns) and of the output clouds
    host_v, host_t = [], []
    for name, v, t in meshes:
        pv = torch.from_numpy(v).pin_memory()
        host_v.append(pv)
        host_t.append(t)
        assert ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0]) >= 0
        assert ctx.update_geometry(name, pv.numpy(), t, frame_affine(0, len(host_v) - 1)) == 0
    n_tris = sum(t.shape[0] for t in host_t)
    commit_mode = capi.COMMIT_REBUILD if args.commit == "rebuild" else capi.COMMIT_AUTO
    names = [m[0] for m in meshes]

    def step_resident(frame):
        for gi, name in enumerate(names):
            ctx.update_transform(name, frame_affine(frame, gi))
        rc = ctx.commit_scene(commit_mode)
        assert rc == 0, ctx.last_error()
        return 0

    def sync_all():
        stream.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident leg (value) ---------------------------------------------------------------------
    frame = 0
    sampler = ClockSampler(local) if rank == 0 else None      # started before the warm-up: sampling, not starting up, during the timed region
    for w in range(args.warmup):
        step_resident(frame * world + rank); rc = ctx.trace_batch(0, n_sens, 0); assert rc == 0, ctx.last_error()
        frame += 1
    if sampler:
        sampler.wait_first()
    sync_all()
    ctx.reset_counters()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    t_wall0 = time.time()
    e_begin, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e_begin.record(stream)
    for k in range(args.steps):
        ev[k][0].record(stream)
        step_resident(frame * world + rank)
        ev[k][1].record(stream)
        rc = ctx.trace_batch(0, n_sens, 0)
        ev[k][2].record(stream)
        frame += 1
    e_end.record(stream)
    sync_all()
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    assert rc == 0, ctx.last_error()
    ms_total = e_begin.elapsed_time(e_end)
    launches = int(ctx.stats().kernel_launches)
    ms_commit = float(np.mean([ev[k][0].elapsed_time(ev[k][1]) for k in range(args.steps)]))
    ms_trace = float(np.mean([ev[k][1].elapsed_time(ev[k][2]) for k in range(args.steps)]))
    if world > 1:
        tt = torch.tensor([ms_total, ms_trace, ms_commit], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total, ms_trace, ms_commit = [float(x) for x in tt.tolist()]
    rays_per_step_rank = n_sens * R
    value = world * rays_per_step_rank * args.steps / (ms_total * 1e-3) / 1e6

    # hit fraction (for the byte model) from the last frame
    rc, counts = ctx.fetch_batch(0, n_sens, [0] * n_sens, [0] * n_sens, 0)
    hits = int(counts.sum())
    hit_fraction = hits / rays_per_step_rank

    # stage breakdown of one more frame with per-stage events (outside the timed region)
    ctx.set_option(capi.OPT_STAGE_TIMING, 1)
    step_resident(frame * world + rank); ctx.trace_batch(0, n_sens, 0)
    st = ctx.stats()
    ctx.set_option(capi.OPT_STAGE_TIMING, 0)
    stages = {k: round(float(getattr(st, "ms_" + k)), 4) for k in
              ("transform", "morton", "sort", "hierarchy", "fit", "toplevel", "trace", "compact")}
    # the trace stage = k_tile_entries (per-tile entry lists, rebuilt with every BVH) + k_trace
    stages["tile_lists"] = round(float(st.ms_tile_lists), 4)
    stages["k_trace"] = round(float(st.ms_trace) - float(st.ms_tile_lists), 4)

    # ---- other splits of configs[3] (extra keys; the headline above shards FRAMES) -----------------------------------------
    def timed_steps(fn, k_steps, warm=2):
        for w_ in range(warm):
            fn(10_000 + w_)
        sync_all()
        a, bb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for kk in range(k_steps):
            fn(20_000 + kk)
        bb.record(stream)
        sync_all()
        ms_ = a.elapsed_time(bb) / k_steps
        if world > 1:
            tt_ = torch.tensor([ms_], device="cuda", dtype=torch.float64)
            dist.all_reduce(tt_, op=dist.ReduceOp.MAX)
            ms_ = float(tt_.item())
        return ms_

    # device-major (SURVEY 8e for C4): ONE frame at a time on all GPUs, the 64 devices sharded over the ranks; every
    # rank pays the full rebuild, so the frame latency is ~ commit + trace / N (the replicated build is the Amdahl term)
    lo_s, hi_s = n_sens * rank // world, n_sens * (rank + 1) // world

    def step_device_major(f):
        step_resident(f)                                    # the SAME frame on every rank
        assert ctx.trace_batch(lo_s, hi_s - lo_s, 0) == 0, ctx.last_error()
    ms_dm = timed_steps(step_device_major, args.steps)
    c4_device_major = {"what": "one frame at a time, 64 devices sharded over the ranks, every rank rebuilds the whole BVH",
                       "ms_per_frame": ms_dm, "mrays_per_s": n_sens * R / (ms_dm * 1e-3) / 1e6, "frames_per_s": 1e3 / ms_dm,
                       "devices_per_gpu": hi_s - lo_s}

    # moving sensors (vehicle-mounted LiDARs): every device's pose changes every frame, so the octant queues, tile
    # intervals, origin chains and tile lists are rebuilt per frame instead of being cached per sensor table
    def step_moving(f):
        for k_, (q_, t_) in enumerate(poses):
            tq = np.array(t_, np.float32); tq[0] += np.float32(0.01 * (f % 100)); tq[1] -= np.float32(0.007 * (f % 100))
            ctx.set_sensor_pose(q_, tq, index=k_)
        step_resident(f * world + rank)
        assert ctx.trace_batch(0, n_sens, 0) == 0, ctx.last_error()
    ms_mv = timed_steps(step_moving, args.steps)
    for k_, (q_, t_) in enumerate(poses):
        ctx.set_sensor_pose(q_, t_, index=k_)
    c4_moving = {"what": "frames sharded over ranks as in the headline, but all 64 sensor poses change every frame "
                         "(lsb200_set_sensor_pose): queues / intervals / chains / lists rebuilt per frame",
                 "ms_per_step": ms_mv, "mrays_per_s": world * rays_per_step_rank / (ms_mv * 1e-3) / 1e6}

    # ---- optional gather of the per-GPU clouds over NVLink (N > 1): shard.gather_clouds on NCCL with the real clouds ----
    gather = None
    if world > 1:
        import ctypes as C
        from lidarshooter_b200 import shard
        rc = ctx.trace_batch(0, n_sens, capi.TRACE_XYZI16)
        assert rc == 0, ctx.last_error()
        rc, counts = ctx.fetch_batch(0, n_sens, [0] * n_sens, [0] * n_sens, capi.TRACE_XYZI16)
        parts = []
        for k in range(n_sens):
            dp = C.c_void_p()
            capi.lib().lsb200_device_results(ctx._h, k, C.byref(dp), None, None, None)
            parts.append(torch.as_tensor(shard.DeviceArray(dp.value, int(counts[k]) * 16), device="cuda").view(-1, 16))
        stream.synchronize()
        cloud = torch.cat(parts)                           # this rank's merged cloud (device memory)
        for _ in range(2):
            shard.gather_clouds(cloud, dst=0)
        sync_all()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        greps = 5
        g0.record()
        for _ in range(greps):
            merged = shard.gather_clouds(cloud, dst=0)
        g1.record()
        torch.cuda.synchronize()
        gms = g0.elapsed_time(g1) / greps
        tt = torch.tensor([gms, float(cloud.shape[0])], device="cuda", dtype=torch.float64)
        mx = tt.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tt.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        total_pts = int(sm[1].item())
        inbound = (total_pts - int(cloud.shape[0])) * 16 if rank == 0 else 0
        if rank == 0:
            assert merged.shape[0] == total_pts
            gather = {"what": "shard.gather_clouds: every rank's merged 16-byte XYZI cloud of one frame to rank 0 "
                              "(all_gather of counts + batched variable-length isend/irecv, NCCL over NVLink / NVSwitch)",
                      "ms": float(mx[0].item()), "points_total": total_pts, "bytes_into_rank0": int(inbound),
                      "gbs_into_rank0": inbound / (float(mx[0].item()) * 1e-3) / 1e9, "timed": "CUDA events, max over ranks, 5 reps"}
        del merged, cloud, parts

    # ---- parity of the benchmarked configuration: per-ray results of one frame, compared with the oracle below ----
    parity_frame = 3
    parity_sensors = (0,) if args.small else (0, 27, 63)
    gpu_results = None
    if rank == 0 and not args.no_parity:
        step_resident(parity_frame)
        gpu_results = {}
        for si in parity_sensors:
            rc_, t_, tri_ = ctx.trace_raw(si)
            assert rc_ == 0, ctx.last_error()
            gpu_results[si] = (t_, tri_)

    # ---- end-to-end leg: host buffers in, host point clouds out --------------------------------------------
    # The public call for a batch of frames is lsb200_run_frames: per frame it uploads every mesh's vertices from
    # pinned host memory, applies the frame's transforms, rebuilds, traces all devices and copies the hit counts and
    # the point clouds into the caller's pinned host buffers; the D2H of frame f overlaps frame f+1's compute.
    # Headline record = the 16-byte XYZI point north_star names; the 32-byte PointCloud2 record (the XT-32
    # message layout, SRC/XYZIRBytes.cpp:24-40) is measured the same way and reported beside it.
    # All pinned host memory of this leg comes from the library (lsb200_alloc_host): allocated and first-touched while
    # the thread is bound to the CPUs of the NUMA node this rank's GPU hangs off, so that with 8 ranks the D2H streams
    # spread over the sockets' memory controllers (tools/d2h_probe.cu measures plain vs placed buffers).
    n_sets = 2
    out_sets = [ctx.alloc_host(n_sens * R * 32) for _ in range(n_sets)]
    out_ptrs = [[int(buf.ctypes.data) + k * R * 32 for k in range(n_sens)] for buf in out_sets]
    caps = [R] * n_sens
    pinned_v = []
    for v in host_v:
        pv = ctx.alloc_host(v.numel() * 4)
        pv[:] = v.numpy().view(np.uint8).reshape(-1)
        pinned_v.append(pv)
    h2d = sum(v.numel() * 4 for v in host_v) + 48 * len(names)
    vert_ptrs = [int(pv.ctypes.data) for pv in pinned_v]

    def run_e2e(first_frame, n_frames, flags):
        A = np.stack([np.stack([frame_affine((first_frame + f) * world + rank, gi) for gi in range(len(names))])
                      for f in range(n_frames)])
        ptrs = [out_ptrs[f % n_sets] for f in range(n_frames)]
        rc_, cnt = ctx.run_frames(A, ptrs, caps, first=0, count=n_sens, commit_mode=commit_mode, flags=flags,
                                  verts=vert_ptrs)
        assert rc_ == 0, ctx.last_error()
        return cnt

    def reduce_max_ms(ms):
        if world > 1:
            tt = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            return float(tt.item())
        return ms

    e2e = {}
    legs = (("xyzi16", capi.TRACE_XYZI16, 16), ("pointcloud2_32", 0, 32), ("range_image", capi.TRACE_RANGE_IMAGE, 4))
    for label, flags, rec in legs:
        run_e2e(frame, 2, flags); frame += 2                       # warm-up (allocates the second output slot)
        sync_all()
        t0 = time.perf_counter()
        cnt = run_e2e(frame, args.steps, flags); frame += args.steps
        sync_all()
        ms = reduce_max_ms((time.perf_counter() - t0) * 1e3)       # blocking call: wall clock covers every copy
        pts_per_step = int(cnt.sum()) // args.steps
        # the clouds really arrived: last frame's first record of sensor 0 is a finite point with intensity 64
        last = out_sets[(args.steps - 1) % n_sets][:32].view(np.float32)
        if rec == 4:
            assert np.isfinite(out_sets[(args.steps - 1) % n_sets][:R * 4].view(np.float32)).any()
        else:
            assert np.isfinite(last[:3]).all() and last[3 if rec == 16 else 4] == 64.0
        d2h = int(pts_per_step * rec + (4 * n_sens if rec != 4 else 0))
        # host ceiling: the SAME copies (64 clouds of the last frame, device -> the same pinned buffers) with nothing else
        # running, all ranks at once -- what this box's PCIe / host memory path gives this copy pattern
        ctx.trace_batch(0, n_sens, flags)
        sync_all()
        t0 = time.perf_counter()
        ceil_reps = 5
        for _ in range(ceil_reps):
            rc_, _cnt = ctx.fetch_batch(0, n_sens, out_ptrs[0], caps, flags)
            assert rc_ == 0, ctx.last_error()
        ms_copy = reduce_max_ms((time.perf_counter() - t0) * 1e3) / ceil_reps
        e2e[label] = {"value": world * rays_per_step_rank * args.steps / (ms * 1e-3) / 1e6, "unit": "Mrays/s",
                      "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": d2h,
                      "ms_per_step": ms / args.steps, "record_bytes": rec,
                      "d2h_gbs_per_gpu": d2h / (ms / args.steps * 1e-3) / 1e9,
                      "host_ceiling_gbs_per_gpu": d2h / (ms_copy * 1e-3) / 1e9,
                      "host_ceiling_ms_per_step": ms_copy,
                      "pinned_memory": f"lsb200_alloc_host (GPU NUMA node {ctx.numa_node()})",
                      "api": "lsb200_run_frames (pipelined: D2H of frame f overlaps frame f+1)"}
    for buf in out_sets + pinned_v:
        ctx.free_host(buf)
    rec = 32

    # ---- roofline ---------------------------------------------------------------------------------------------
    # Dominant kernels = the trace stage (k_tile_entries + k_trace), timed live above with CUDA events on the launching
    # stream (ms_trace, every timed step).  Three figures, each <= ~1 by construction, plus the contract's model:
    #   frac             DRAM bytes the stage really moves (ncu dram__bytes_read + write of THIS build: the capture in
    #                    profiles/ncu_traffic.json is refused unless it is stamped with this tree's source hash)
    #                    / ms_trace / measured HBM peak
    #   algorithmic_frac compulsory bytes of the stage (every node and triangle record once, t + id per ray, one record
    #                    per hit) / ms_trace / peak
    #   step_floor_frac  compulsory bytes of the WHOLE step (rebuild 314 B/triangle, SURVEY 8d, + the stage's) /
    #                    ms_per_step / peak
    #   contract_frac    SURVEY 8(d)'s per-ray model (a private root-to-leaf path per ray).  Coherent LiDAR rays share
    #                    that path through the tile lists and L1 / L2, so this one exceeds 1: kept for the record only.
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md 6.65 TB/s)"
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from source_hash import source_hash
    src_hash = source_hash()
    traffic, ncu_k, ncu_note = None, {}, "no profiles/ncu_traffic.json"
    try:
        nt = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if nt.get("source_hash") == src_hash:
            ncu_k = nt["kernels"]
            traffic = float(ncu_k["k_trace"]["dram_bytes"]) + float(ncu_k.get("k_tile_entries", {}).get("dram_bytes", 0.0))
            ncu_note = "profiles/ncu_traffic.json (stamped with this build's source hash)"
        else:
            ncu_note = (f"profiles/ncu_traffic.json REFUSED: captured from source {nt.get('source_hash')}, "
                        f"this build is {src_hash}")
    except Exception:
        pass
    stage_s = ms_trace * 1e-3
    stage_alg_bytes = (64.0 + 48.0 + 4.0) * n_tris + rays_per_step_rank * (8.0 + rec * hit_fraction)
    build_alg_bytes = (314.0 if args.commit == "rebuild" else 214.0) * n_tris
    contract_bytes = rays_per_step_rank * b_ray(n_tris, hit_fraction)
    roofline = {
        "kernel": "trace stage = k_tile_entries + k_trace<6> (one launch each per step)", "bound": "hbm",
        "achieved": (traffic / stage_s / 1e9) if traffic else stage_alg_bytes / stage_s / 1e9,
        "peak": peak, "unit": "GB/s", "peak_source": peak_src,
        "traffic": traffic, "traffic_source": ncu_note, "source_hash": src_hash,
        "ms_per_launch": ms_trace, "rays_per_launch": rays_per_step_rank,
        "algorithmic_bytes": stage_alg_bytes, "algorithmic_frac": stage_alg_bytes / stage_s / 1e9 / peak,
        "step_floor_bytes": build_alg_bytes + stage_alg_bytes,
        "step_floor_frac": (build_alg_bytes + stage_alg_bytes) / (ms_total / args.steps * 1e-3) / 1e9 / peak,
        "commit_frac": build_alg_bytes / (ms_commit * 1e-3) / 1e9 / peak,
        "contract_bytes_per_ray": b_ray(n_tris, hit_fraction), "contract_achieved": contract_bytes / stage_s / 1e9,
        "contract_frac": contract_bytes / stage_s / 1e9 / peak,
    }
    roofline["frac"] = roofline["achieved"] / peak
    roofline["frac_is"] = ("ncu DRAM traffic of this build / live stage time / measured peak" if traffic else
                           "algorithmic_frac (no valid ncu capture of this build)")
    if "k_trace" in ncu_k:
        kt = ncu_k["k_trace"]
        roofline["actual_bound"] = {
            "what": "k_trace is latency / issue bound, not bandwidth bound (ncu of this build)",
            "issue_slots_busy_pct": kt.get("issue_slots_busy_pct"), "warp_exec_eff_threads_of_32": kt.get("warp_exec_eff_threads"),
            "l1_hit_pct": kt.get("l1_hit_pct"), "l2_hit_pct": kt.get("l2_hit_pct"), "warps_active_pct": kt.get("warps_active_pct"),
            "registers": kt.get("registers"), "ncu_duration_ms": (kt.get("duration_ns") or 0.0) * 1e-6}
        roofline["per_kernel"] = {k: {"dram_bytes": v.get("dram_bytes"), "ncu_duration_ms": (v.get("duration_ns") or 0.0) * 1e-6,
                                      "dram_frac_of_peak": (v.get("dram_bytes", 0.0) / max(v.get("duration_ns", 1.0), 1.0)) / peak}
                                  for k, v in ncu_k.items()}

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args.small), "n_triangles": n_tris, "devices_per_gpu": n_sens,
                       "rays_per_device": R, "rays_per_step_per_gpu": rays_per_step_rank, "frames_per_step": world,
                       "commit": args.commit, "sharding": "trajectory frames over ranks, scene replicated per GPU",
                       "l2": "inputs larger than L2 (BVH + triangles 1.1 GB, outputs 1.2 GB per frame); no flush needed",
                       "point_record_bytes": rec, "hit_fraction": round(hit_fraction, 4)},
            "scans_per_s": world * n_sens * args.steps / (ms_total * 1e-3),
            "trace_only_mrays_per_s": world * rays_per_step_rank / (ms_trace * 1e-3) / 1e6,
            "ms_commit": ms_commit, "ms_trace": ms_trace, "stages_ms": stages,
            "clocks": clocks,
            "e2e": e2e["xyzi16"], "e2e_pointcloud2_32": e2e["pointcloud2_32"], "e2e_range_image": e2e["range_image"],
            "gpu_launches": launches,
            "gather": gather, "c4_device_major": c4_device_major, "c4_moving_sensors": c4_moving,
            "roofline": roofline,
        }
    ctx.close()
    if rank == 0 and not args.no_xt32:
        line["xt32_scan"] = xt32_scan_leg(local, cpu=(world == 1 and not args.no_cpu_baseline))
    if not args.no_other_configs:
        oc = other_configs_leg(local, rank, world)
        if rank == 0:
            line["other_configs"] = oc
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        want_cpu = world == 1 and not args.no_cpu_baseline
        c = None
        if want_cpu or gpu_results is not None:
            log("[bench] CPU oracle on the host cores: parity of the benchmarked frame + bounded timing sample ...")
            try:      # all the host threads the process may use -- torchrun exports OMP_NUM_THREADS=1 to its workers
                n_thr = len(os.sched_getaffinity(0))
            except AttributeError:
                n_thr = os.cpu_count() or 1
            c = cpu_trace_sample(meshes, poses, table, n_threads=n_thr, sensors=parity_sensors if gpu_results else (0,),
                                 affines=[frame_affine(parity_frame, gi) for gi in range(len(names))], gpu_results=gpu_results)
            line["parity_check"] = c["parity"]
        if want_cpu:
            packet = c["mrays_packet"] is not None
            line["cpu_baseline"] = {
                # the faster of the two CPU forms is the baseline: 8-ray AVX2 packets when the host has AVX2 (closer to
                # the reference's rtcIntersect16), else the scalar walk; both are reported
                "value": max(c["mrays_packet"], c["mrays_trace"]) if packet else c["mrays_trace"],
                "unit": "Mrays/s", "cores": c["cores"], "kind": "port",
                "form": "8-ray AVX2 packets" if packet else "scalar rays",
                "scalar_value": c["mrays_trace"], "packet_value": c["mrays_packet"],
                "packet_same_as_scalar": c["packet_same_as_scalar"],
                "sample": (f"closest-hit trace of {len(parity_sensors) if gpu_results else 1} of the {n_sens} devices ({c['rays']} rays) against the same "
                           f"{c['n_tris']}-triangle scene with the oracle's binned-SAH CPU BVH, all host threads, traced twice: one ray at a "
                           f"time and as 8-ray AVX2 packets (bit-identical results); "
                           f"trace only -- CPU transform {c['t_transform']:.2f} s and BVH build {c['t_build']:.2f} s "
                           f"are NOT charged (the GPU value pays its rebuild every frame).  Embree 3.13.4 itself: not installed, not installable"),
                "t_build_s": c["t_build"], "t_trace_s": c["t_trace"], "t_packet_s": c["t_packet"]}
        else:
            line["cpu_baseline"] = None
        emit_result(line)
        pc = line.get("parity_check")
        if pc and (pc["mismatch_mask"] or pc["mismatch_tri"] or pc["mismatch_t"] or pc["grazing_unexplained"]):
            raise SystemExit(f"bench.py: PARITY FAILURE on the benchmarked configuration: {pc}")


if __name__ == "__main__":
    main()
