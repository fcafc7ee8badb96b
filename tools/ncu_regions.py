#!/usr/bin/env python
"""Samples / instructions per address region of an ncu source-page csv.
usage: ncu_regions.py <sass.csv> name:lo:hi [name:lo:hi ...]   (hex offsets)"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
regions = [(a.split(':')[0], int(a.split(':')[1], 16), int(a.split(':')[2], 16)) for a in sys.argv[2:]]
i = 0
while not rows[i] or rows[i][0] != 'Kernel Name': i += 1
hdr = rows[i + 1]
ca, cs, cw, ce = hdr.index('Address'), hdr.index('Source'), hdr.index('Warp Stall Sampling (All Samples)'), hdr.index('Instructions Executed')
stall_cols = [(j, h[6:]) for j, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
base = None
agg = {r[0]: [0, 0, collections.Counter()] for r in regions}
for r in rows[i + 2:]:
    if not r or r[0] == 'Kernel Name': break
    a = int(r[ca], 16) if r[ca].startswith('0x') else int(r[ca])
    if base is None: base = a
    off = a - base
    for name, lo, hi in regions:
        if lo <= off < hi:
            agg[name][0] += int(r[cw] or 0); agg[name][1] += int(r[ce] or 0)
            for j, h in stall_cols:
                if r[j] and int(r[j]): agg[name][2][h] += int(r[j])
for name, _, _ in regions:
    s, e, c = agg[name]
    print(f'{name:6s} samples {s:5d} inst/tile {e/3008:8.1f}  ', ', '.join(f'{k}:{v}' for k, v in c.most_common(7)))
