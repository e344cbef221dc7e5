#!/usr/bin/env python
"""Cycles per phase of k_onesweep (thread 0 of every CTA, averaged over tiles and passes) on n random 30-bit pairs.
Needs a library built with -DLSB_SORT_PROF (LSB200_LIB=...); evidence tool, not the product path."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lidarshooter_b200 import capi  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
rng = np.random.default_rng(1)
keys = rng.integers(0, 1 << 30, n, dtype=np.uint32)
vals = np.arange(n, dtype=np.uint32)
ctx = capi.Context(0)
L = capi.lib()
out = (C.c_ulonglong * 8)()
ctx.debug_sort_pairs(keys, vals)
L.lsb200_debug_sort_prof(out)
ctx.debug_sort_pairs(keys, vals)
L.lsb200_debug_sort_prof(out)
tiles = -(-n // 8192) * 4
names = ["ticket+zero+barrier+load issue", "load wait + ranking (warp 0)", "barrier after ranking", "scan + look-back (digit 0)",
         "block scan + barrier", "reorder (thread 0)", "barrier after reorder", "write-out issue"]
tot = sum(out)
for k, nm in enumerate(names):
    print(f"{nm:34s} {out[k] / tiles:9.0f} cycles / tile  {100 * out[k] / tot:5.1f} %")
print(f"{'total':34s} {tot / tiles:9.0f} cycles / tile")
