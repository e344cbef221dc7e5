#!/usr/bin/env python
"""Region-level split (instructions, stall samples, active threads) of k_trace<6> from an ncu SASS export joined with
the cubin's line table (tools/sass_lines.py).  usage: sass_regions.py source.csv [kernel substring]"""
import csv
import os
import sys
from collections import defaultdict
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sass_lines as sl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def regions():
    """Line ranges of trace.cuh by function, found from the source so they survive edits."""
    src = open(os.path.join(ROOT, 'lidarshooter_b200', 'csrc', 'trace.cuh')).read().splitlines()
    marks = [('intersect_tri', 'tri test'), ('void test_leaf', 'leaf loop'), ('void slab_nf', 'slab_nf'),
             ('void trace_ray(', 'trace_ray (modes 1-4)'), ('void trace_ray_ww', 'ww: set-up + entry loop'),
             ('    while (cur != kDone) {', 'ww: node loop'), ('        if (cur < 0) {', 'ww: leaf + pop'),
             ('void trace_ray_lz', 'lz: lazy-list traversal (list tests, node loop, pops)'), ('void trace_ray_il', 'il: single-loop traversal'),
             ('void brute_ray', 'brute'), ('void store_point', 'store/hit point'), ('void ray_from_tables', 'make_ray'),
             ('int ray_octant', 'unit descriptor'), ('void process_unit', 'process_unit'), ('void cp_async16', 'run_phase'),
             ('void k_trace(', 'kernel'), ('void __launch_bounds__(256) k_classify_units', 'other kernels')]
    out = []
    for i, ln in enumerate(src, 1):
        for m, name in marks:
            if m in ln and not any(n == name for _, n in out):
                out.append((i, name))
    return sorted(out)


def main():
    srcf = sys.argv[1]
    kernel = sys.argv[2] if len(sys.argv) > 2 else 'k_traceILi6ELb0'
    table = sl.line_table(os.path.join(ROOT, 'lidarshooter_b200', 'liblsb200.so'), kernel)
    rg = regions()
    rows = list(csv.reader(open(srcf)))
    hdr = next(r for r in rows if r and r[0] == 'Address')
    ix = {h: i for i, h in enumerate(hdr)}
    ie, te, ss = ix['Instructions Executed'], ix['Thread Instructions Executed'], ix['Warp Stall Sampling (All Samples)']
    data = [r for r in rows if len(r) > te and r[ie].isdigit()]
    base = min(int(r[0], 16) for r in data)
    seen, agg = set(), defaultdict(lambda: [0, 0, 0])
    for r in data:
        off = int(r[0], 16) - base
        if off in seen:
            continue
        seen.add(off)
        key = table.get(off)
        name = 'other: ' + (key[0] if key else '?')
        if key and key[0] == 'trace.cuh':
            name = [n for l, n in rg if l <= key[1]][-1] if key[1] >= rg[0][0] else 'header'
        g = agg[name]; g[0] += int(r[ie]); g[1] += int(r[te]); g[2] += int(r[ss])
    tot = sum(a[0] for a in agg.values()); tots = max(sum(a[2] for a in agg.values()), 1)
    print(f'{kernel}: {tot} warp instructions, {tots} stall samples')
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f'{k:36s} inst {a[0] / tot * 100:5.1f}%  stall {a[2] / tots * 100:5.1f}%  threads {a[1] / max(a[0], 1):5.1f}')


if __name__ == '__main__':
    main()
