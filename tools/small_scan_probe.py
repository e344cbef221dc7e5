#!/usr/bin/env python
"""Probe of the small-scene path: a few XT-32 scans of ground + ben (prints the fused build kernel's phase clocks when
the library was built with -DLSB_SMALL_CLK)."""
import os
import sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import GOLDEN, xt32_sensor_args
from lidarshooter_b200 import capi
z = np.load(os.path.join(GOLDEN, "meshes.npz"))
ctx = capi.Context(0, lab=True)
ctx.set_sensor(*xt32_sensor_args("0000"))
ident = [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0]
for name in ("ground", "ben"):
    v, t = z[name + "_v"], z[name + "_t"]
    ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0])
    ctx.update_geometry(name, v, t, ident)
ctx.set_option(capi.OPT_STAGE_TIMING, 1)
for mode, min_units in ((1, 4096), (2, 4096), (4, 4096), (0, 4096), (6, 4096), (6, 0), (7, 0)):
    ctx.set_option(capi.OPT_TRAVERSAL, mode)
    ctx.set_option(capi.OPT_TILE_MIN_UNITS, min_units)       # 0: per-tile entry lists even for the 150 tiles of an XT-32
    ts = []
    for k in range(8):
        assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0
        ctx.trace_device(0, 0)
        ts.append(ctx.stats().ms_trace * 1e3)
    print("mode", mode, "tile lists" if min_units == 0 else "no tile lists", "trace us", " ".join("%.1f" % t for t in ts), flush=True)
ctx.set_option(capi.OPT_TRAVERSAL, 6)
ctx.set_option(capi.OPT_TILE_MIN_UNITS, 4096)
for k in range(3):
    assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0
    rc, pts = ctx.trace_scene(0, 0)
    ctx.synchronize()
    st = ctx.stats()
    print("points", pts.shape[0], "build us %.1f  trace us %.1f  compact us %.1f" % (st.ms_fit * 1e3, st.ms_trace * 1e3, st.ms_compact * 1e3), flush=True)
ctx.close()
