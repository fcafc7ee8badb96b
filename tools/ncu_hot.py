#!/usr/bin/env python
"""Hot SASS instructions of an ncu source-page csv: sync points and anything above a sample threshold.
usage: ncu_hot.py <sass.csv> [min samples] [units per launch]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = int(sys.argv[2]) if len(sys.argv) > 2 else 60
U = float(sys.argv[3]) if len(sys.argv) > 3 else 3008.0
i = 0
while not rows[i] or rows[i][0] != 'Kernel Name': i += 1
hdr = rows[i + 1]
ca, cs, cw, ce = hdr.index('Address'), hdr.index('Source'), hdr.index('Warp Stall Sampling (All Samples)'), hdr.index('Instructions Executed')
stall_cols = [(j, h) for j, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
base = None; tot = 0; out = []
for r in rows[i + 2:]:
    if not r or r[0] == 'Kernel Name': break
    a = int(r[ca], 16) if r[ca].startswith('0x') else int(r[ca])
    if base is None: base = a
    s = int(r[cw] or 0); ex = int(r[ce] or 0); tot += s
    if 'SYNCS' in r[cs] or 'BAR' in r[cs] or s > thr:
        kinds = sorted(((int(r[j] or 0), h[6:]) for j, h in stall_cols), reverse=True)[:2]
        out.append((a - base, r[cs][:64], s, ex, ' '.join(f'{k}:{v}' for v, k in kinds if v)))
print('total samples', tot)
for o in out: print(f'{o[0]:6x} {o[1]:64s} smp {o[2]:5d} exec {o[3]/U:8.1f}/u  {o[4]}')
