#!/bin/bash
# usage: tools/prof_logmel.sh <tag>   -- ncu full capture of the log-mel kernel + per-region summary
TAG=$1
/usr/local/graft/bin/gpurun --timeout 900 -- "python bench.py --no-cpu-baseline --steps 8 --warmup 3 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:logmel_kernel -s 4 -c 1 -o gpurun_out/${TAG}_logmel python bench.py --no-cpu-baseline --steps 8 --warmup 3 > gpurun_out/ncu_${TAG}.log 2>&1; tail -1 gpurun_out/ncu_${TAG}.log" 2>&1 | tail -2
ncu -i gpurun_out/${TAG}_logmel.ncu-rep --page raw --csv > gpurun_out/${TAG}_logmel_raw.csv 2>/dev/null
python tools/ncu_keys.py gpurun_out/${TAG}_logmel_raw.csv | sed -n 4,40p
ncu -i gpurun_out/${TAG}_logmel.ncu-rep --page source --csv --print-source sass > gpurun_out/${TAG}_logmel_sass.csv 2>/dev/null
python tools/ncu_lines.py gpurun_out/${TAG}_logmel_sass.csv whisper_rs_b200/lib/libwrs_b200.so logmel_kernelILb1 2000 > gpurun_out/${TAG}_lines.txt
python - <<PY
import re,collections
rows=[]
for ln in open('gpurun_out/${TAG}_lines.txt').read().splitlines()[1:]:
    m=re.match(r"\s*([\d.]+)%\s+inst\s+(\d+)\s+\('(.*?)', (\d+)\)\s+\[(.*)\]",ln)
    if m: rows.append((m.group(3),int(m.group(4)),int(m.group(2)),float(m.group(1)),m.group(5)))
print('top stall lines:')
for f,l,n,p,st in sorted(rows,key=lambda r:-r[3])[:12]: print(f'  {p:5.1f}%  inst {n/3008.:8.1f}/slice  {f}:{l}  {st}')
print('top inst lines:')
for f,l,n,p,st in sorted(rows,key=lambda r:-r[2])[:16]: print(f'  {p:5.1f}%  inst {n/3008.:8.1f}/slice  {f}:{l}  {st}')
print('total inst/slice', sum(r[2] for r in rows)/3008.)
PY
