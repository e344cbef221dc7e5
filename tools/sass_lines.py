#!/usr/bin/env python
"""Per-source-line profile of one kernel: joins an `ncu --page source --csv` export (SASS view: executed warp
instructions, stall samples per address) with the line table of the cubin (`nvdisasm -g`), so that instructions
and stall samples are attributed to lines of trace.cuh / build.cuh.  The .so must be the build that was profiled.
usage: sass_lines.py source.csv <mangled kernel name substring> [liblsb200.so] [n_lines]"""
import csv
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def line_table(so, kernel):
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(['cuobjdump', '-xelf', 'all', so], cwd=d, check=True, capture_output=True)
        cubin = [os.path.join(d, f) for f in os.listdir(d) if f.endswith('.cubin')][0]
        txt = subprocess.run(['nvdisasm', '-g', '-c', cubin], capture_output=True, text=True).stdout
    table, cur, inside = {}, None, False
    for ln in txt.splitlines():
        if ln.startswith('\t.section\t.text.'):
            inside = kernel in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            # inlined-at chains print several File lines; the LAST one before an instruction is the innermost
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r'\s*/\*([0-9a-f]{4,})\*/', ln)
        if m:
            table[int(m.group(1), 16)] = cur
    return table


def main():
    src, kernel = sys.argv[1], sys.argv[2]
    so = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, 'lidarshooter_b200', 'liblsb200.so')
    ntop = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    table = line_table(os.path.abspath(so), kernel)
    rows = list(csv.reader(open(src)))
    hdr = next(r for r in rows if r and r[0] == 'Address')
    ix = {h: i for i, h in enumerate(hdr)}
    ie, te, ss = ix['Instructions Executed'], ix['Thread Instructions Executed'], ix['Warp Stall Sampling (All Samples)']
    data = [r for r in rows if len(r) > te and r[ie].isdigit()]
    base = min(int(r[0], 16) for r in data)
    agg = defaultdict(lambda: [0, 0, 0])
    seen = set()
    for r in data:
        off = int(r[0], 16) - base
        if off in seen:          # several launches in one export: count the first
            continue
        seen.add(off)
        a = agg[table.get(off)]
        a[0] += int(r[ie]); a[1] += int(r[te]); a[2] += int(r[ss])
    tot = sum(a[0] for a in agg.values()); tots = max(sum(a[2] for a in agg.values()), 1)
    print(f'kernel {kernel}: {tot} warp instructions, {tots} stall samples, {len(table)} SASS lines mapped')
    cache = {}
    def text(key):
        if not key:
            return '?'
        f, n = key
        for d in ('lidarshooter_b200/csrc',):
            p = os.path.join(ROOT, d, f)
            if os.path.exists(p):
                if p not in cache:
                    cache[p] = open(p).read().splitlines()
                return cache[p][n - 1].strip()[:90] if n - 1 < len(cache[p]) else ''
        return ''
    print('  inst%  stall%  thr  line')
    for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0] - kv[1][2] * tot / tots)[:ntop]:
        print(f'{a[0] / tot * 100:6.1f} {a[2] / tots * 100:6.1f} {a[1] / max(a[0], 1):5.1f}  {key[0] + ":" + str(key[1]) if key else "?":22s} {text(key)}')


if __name__ == '__main__':
    main()
