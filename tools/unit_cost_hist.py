import sys, os
sys.path.insert(0, '/root/repo')
import numpy as np
from bench import build_workload, frame_affine
from lidarshooter_b200 import capi
meshes, poses, table = build_workload(False)
elev, b, e, H = table
ctx = capi.Context(0, capi.FRAME_WORLD, lab=True)
for k, (q, t) in enumerate(poses):
    ctx.set_sensor(elev, b, e, H, q, t, index=k)
for gi, (name, v, t) in enumerate(meshes):
    ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0]); ctx.update_geometry(name, v, t, frame_affine(0, gi))
assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0
ctx.set_option(capi.OPT_TRACE_STATS, 1)
assert ctx.trace_batch(0, len(poses), 0) == 0
c = ctx.trace_counters()
tot = sum(c['cyc_sum'])
print('bin  units   %units  %cycles  nodes/ray  list')
for bn in range(24):
    u = c['cyc_units'][bn]
    if u: print(bn, u, round(100*u/sum(c['cyc_units']),2), round(100*c['cyc_sum'][bn]/tot,2), round(c['cyc_nodes'][bn]/u/32,2), round(c['cyc_list'][bn]/u,1))
hc, hb = c['heavy_cycles_by_class'], c['heavy_cycles_by_band']
print('heavy-unit cycles by octant class (0..7, 8 = mixed), % of all unit cycles:', [round(100 * x / tot, 1) for x in hc])
print('heavy-unit cycles by elevation band (32 bands, +15 deg first .. -25 deg last), %:', [round(100 * x / tot, 1) for x in hb])
