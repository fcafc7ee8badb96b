#!/bin/bash
# usage: tools/prof_decode.sh <tag>   -- ncu full capture of the decode step kernel + per-line summary
TAG=$1
/usr/local/graft/bin/gpurun --timeout 900 -- "python bench.py --no-cpu-baseline --steps 8 --warmup 3 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:${KERN:-decode_scan_kernel} -s 6 -c 1 -o gpurun_out/${TAG}_decode python bench.py --no-cpu-baseline --steps 8 --warmup 3 > gpurun_out/ncu_${TAG}.log 2>&1; tail -1 gpurun_out/ncu_${TAG}.log | cut -c1-200" 2>&1 | tail -2
ncu -i gpurun_out/${TAG}_decode.ncu-rep --page raw --csv > gpurun_out/${TAG}_decode_raw.csv 2>/dev/null
python tools/ncu_keys.py gpurun_out/${TAG}_decode_raw.csv | sed -n 4,40p
ncu -i gpurun_out/${TAG}_decode.ncu-rep --page source --csv --print-source sass > gpurun_out/${TAG}_decode_sass.csv 2>/dev/null
python tools/ncu_lines.py gpurun_out/${TAG}_decode_sass.csv whisper_rs_b200/lib/libwrs_b200.so ${KERN:-decode_scan_kernel} 2000 > gpurun_out/${TAG}_lines.txt
python - <<PY
import re
rows=[]
for ln in open('gpurun_out/${TAG}_lines.txt').read().splitlines()[1:]:
    m=re.match(r"\s*([\d.]+)%\s+inst\s+(\d+)\s+\('(.*?)', (\d+)\)\s+\[(.*)\]",ln)
    if m: rows.append((m.group(3).split('/')[-1],int(m.group(4)),int(m.group(2)),float(m.group(1)),m.group(5)))
U=9440.
print('top stall lines:')
for f,l,n,p,st in sorted(rows,key=lambda r:-r[3])[:14]: print(f'  {p:5.1f}%  inst {n/U:7.1f}/unit  {f}:{l}  {st[:80]}')
print('top inst lines:')
for f,l,n,p,st in sorted(rows,key=lambda r:-r[2])[:22]: print(f'  {p:5.1f}%  inst {n/U:7.1f}/unit  {f}:{l}  {st[:80]}')
print('total inst/unit', sum(r[2] for r in rows)/U)
PY
