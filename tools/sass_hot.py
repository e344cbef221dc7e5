#!/usr/bin/env python
"""Aggregates an `ncu --page source --csv` export: opcode mix by executed warp instructions and a per-block
profile of where the instructions go.  usage: sass_hot.py source.csv [block]"""
import csv
import sys
from collections import Counter


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    blk = int(sys.argv[2]) if len(sys.argv) > 2 else 200
    hdr = rows[1]
    data = rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    ie, te = ix['Instructions Executed'], ix['Thread Instructions Executed']
    tot = sum(int(r[ie]) for r in data)
    print('total warp instructions', tot, 'SASS lines', len(data))
    c, ct = Counter(), Counter()
    for r in data:
        op = r[ix['Source']].split()
        o = op[0] if not op[0].startswith('@') else op[1]
        o = o.split('.')[0]
        c[o] += int(r[ie])
        ct[o] += int(r[te])
    for o, n in c.most_common(25):
        print(f'{o:10s} {n / tot * 100:5.1f}%  avg threads {ct[o] / max(n, 1):.1f}')
    for b in range(0, len(data), blk):
        s = sum(int(r[ie]) for r in data[b:b + blk])
        print(b, f'{s / tot * 100:.1f}%')


if __name__ == '__main__':
    main()
