#!/bin/bash
# usage: tools/gpu_pj_prof.sh <tag>  -- projection: launch list + ncu --set full of proj_kernel<8> from tools/proj_variants.py run (base variant only)
TAG=$1
B="python tools/proj_variants.py run base"
CMD="$B > gpurun_out/${TAG}_plain.log 2>&1; tail -2 gpurun_out/${TAG}_plain.log; ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $B > gpurun_out/${TAG}_list.log 2>&1; timeout 500 ncu --set full --clock-control none --import-source on -k regex:proj_kernel -s 20 -c 1 -o gpurun_out/${TAG}_proj $B > gpurun_out/${TAG}_ncu.log 2>&1; tail -1 gpurun_out/${TAG}_ncu.log | cut -c1-150"
for try in 1 2 3 4 5 6; do
  /usr/local/graft/bin/gpurun --timeout 1200 -- "$CMD" > /tmp/gpu_pj_$TAG.log 2>&1
  if grep -q "status=transient" /tmp/gpu_pj_$TAG.log; then sleep 100; continue; fi
  break
done
tail -8 /tmp/gpu_pj_$TAG.log
python - <<PY
import csv, collections
rows=list(csv.reader(l for l in open('gpurun_out/${TAG}_launches.csv') if l.startswith('"')))
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value')
d=collections.defaultdict(list)
for r in rows[1:]:
    try: d[r[ki][:60]].append(float(r[vi].replace(',','')))
    except: pass
for k,v in d.items(): print(f"{k:60s} n={len(v):4d} median={sorted(v)[len(v)//2]/1000:8.1f} us")
PY
ncu -i gpurun_out/${TAG}_proj.ncu-rep --page raw --csv > gpurun_out/${TAG}_proj_raw.csv 2>/dev/null
python tools/ncu_keys.py gpurun_out/${TAG}_proj_raw.csv | sed -n 1,45p
