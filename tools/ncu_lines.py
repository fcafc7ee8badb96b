#!/usr/bin/env python
"""Aggregate ncu warp-stall samples per source line.

usage: ncu_lines.py <sass.csv from `ncu -i X.ncu-rep --page source --csv --print-source sass`> <lib.so> <mangled kernel substring> [top N]   (uses the first kernel in the csv)
Maps SASS addresses to file:line through `nvdisasm -g` of the matching cubin (needs -lineinfo)."""
import csv, os, re, subprocess, sys, tempfile, collections

sass_csv, so, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run(['cuobjdump', '-xelf', 'all', os.path.abspath(so)], cwd=tmp, check=True, capture_output=True)
addr2line = {}
for f in os.listdir(tmp):
    out = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(tmp, f)], capture_output=True, text=True).stdout
    cur_fn, cur_line, inside = None, None, False
    for ln in out.splitlines():
        m = re.match(r'\s*\.text\.(\S+):', ln)
        if m:
            inside = kern in m.group(1); continue
        if not inside: continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur_line = (os.path.basename(m.group(1)), int(m.group(2))); continue
        m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);', ln)
        if m:
            addr2line[int(m.group(1), 16)] = (cur_line, m.group(2))
rows = list(csv.reader(open(sass_csv)))
# find the header row of the first kernel whose name matches
agg = collections.Counter(); inst = collections.Counter(); stall_kind = collections.defaultdict(collections.Counter)
i = 0; total = 0; base = None
while i < len(rows):
    if rows[i] and rows[i][0] == 'Kernel Name':
        hdr = rows[i + 1]; i += 2
        ci_addr, ci_s, ci_ex = hdr.index('Address'), hdr.index('Warp Stall Sampling (All Samples)'), hdr.index('Instructions Executed')
        stall_cols = [(j, h) for j, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
        while i < len(rows) and rows[i] and rows[i][0] != 'Kernel Name':
            r = rows[i]; a = int(r[ci_addr], 16) if r[ci_addr].startswith('0x') else int(r[ci_addr])
            if base is None: base = a
            key = addr2line.get(a - base, (None, ''))[0]
            s = int(r[ci_s] or 0); agg[key] += s; total += s; inst[key] += int(r[ci_ex] or 0)
            for j, h in stall_cols:
                if r[j] and int(r[j]): stall_kind[key][h[6:]] += int(r[j])
            i += 1
        break
    i += 1
print(f'total samples {total}; total warp-instructions {sum(inst.values())}')
for key, s in agg.most_common(top):
    kinds = ', '.join(f'{k}:{v}' for k, v in stall_kind[key].most_common(3))
    print(f'{100.0 * s / max(total, 1):5.1f}%  inst {inst[key]:9d}  {key}  [{kinds}]')
