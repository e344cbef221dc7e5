cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
cat > /tmp/xt32.py <<'PY'
import os, sys, numpy as np
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
from conftest import GOLDEN, xt32_sensor_args
from lidarshooter_b200 import capi
z = np.load(os.path.join(GOLDEN, "meshes.npz"))
ctx = capi.Context(0)
ctx.set_sensor(*xt32_sensor_args("0000"))
ident = [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0]
for name in ("ground", "ben"):
    v, t = z[name + "_v"], z[name + "_t"]
    ctx.add_geometry(name, capi.GEOMETRY_TRIANGLE, v.shape[0], t.shape[0]); ctx.update_geometry(name, v, t, ident)
for k in range(8):
    assert ctx.commit_scene(capi.COMMIT_REBUILD) == 0
    ctx.trace_device(0, 0)
ctx.synchronize(); ctx.close()
PY
timeout -s KILL 300 ncu --set full --cache-control none --clock-control none --import-source on -k regex:"k_trace_small|k_build_small" --launch-skip 10 --launch-count 2 -o gpurun_out/l_small python /tmp/xt32.py > gpurun_out/l_ncu.log 2>&1
ncu -i gpurun_out/l_small.ncu-rep --page source --csv > gpurun_out/l_small_src.csv 2>/dev/null
ncu -i gpurun_out/l_small.ncu-rep --page details 2>/dev/null | grep -E "^  [a-z_ ]+.*\(|Duration|Issue Slots Busy|Avg. Active Threads|L2 Hit|L1/TEX Hit|Warp Cycles Per Issued|Executed Instructions  " | head -20
