#!/usr/bin/env python
"""Aggregates an `ncu --page source --csv` export (one or several kernels): opcode mix by executed warp instructions
and stall samples, and the hottest SASS lines.  usage: sass_agg.py source.csv [n_top_lines]"""
import csv
import sys
from collections import Counter


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    hdr = next(r for r in rows if r and r[0] == 'Address')
    ix = {h: i for i, h in enumerate(hdr)}
    ie, te, ss = ix['Instructions Executed'], ix['Thread Instructions Executed'], ix['Warp Stall Sampling (All Samples)']
    data = [r for r in rows if len(r) > te and r[ie].isdigit()]
    tot = sum(int(r[ie]) for r in data)
    tots = max(sum(int(r[ss]) for r in data), 1)
    print('total warp instructions', tot, 'SASS lines', len(data), 'stall samples', tots)
    c, cs, ct = Counter(), Counter(), Counter()
    for r in data:
        op = r[ix['Source']].split()
        o = op[0] if not op[0].startswith('@') else op[1]
        o = '.'.join(o.split('.')[:3]) if o.startswith(('LD', 'ST', 'ATOM', 'RED')) else o.split('.')[0]
        c[o] += int(r[ie]); cs[o] += int(r[ss]); ct[o] += int(r[te])
    for o, n in c.most_common(30):
        print(f'{o:20s} inst {n / tot * 100:5.1f}%  samples {cs[o] / tots * 100:5.1f}%  avg threads {ct[o] / max(n, 1):.1f}')
    print('hottest lines by stall samples: samples, executed, avg threads, SASS')
    for r in sorted(data, key=lambda r: -int(r[ss]))[:ntop]:
        print(r[ss], r[ie], r[ix['Avg. Threads Executed']], r[ix['Source']][:100])


if __name__ == '__main__':
    main()
