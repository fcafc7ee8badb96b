#!/bin/bash
# usage: tools/prof_all.sh <tag>   (e.g. r2_v10)
# One gpurun call: the bench line, the launch list, and one `ncu --set full` capture per kernel of the same command;
# then, here, per-kernel summaries under profiles/ and profiles/r2_traffic.json (what bench.py reads `roofline.traffic` from).
TAG=$1
B="python bench.py --no-cpu-baseline --only decode,kv,project --steps 8 --warmup 3"
CMD="$B > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${TAG}_launches.csv $B > gpurun_out/${TAG}_ncu_list.log 2>&1"
for K in logmel_kernel:4 decode_scan_kernel:6 decode_tail_kernel:6 no_speech_kernel:2 kv_reorder_kernel:2 proj_kernel:6; do
  N=${K%%:*}; S=${K##*:}
  CMD="$CMD; timeout 500 ncu --set full --clock-control none --import-source on -k regex:$N -s $S -c 1 -o gpurun_out/${TAG}_$N $B > gpurun_out/${TAG}_ncu_$N.log 2>&1; tail -1 gpurun_out/${TAG}_ncu_$N.log | cut -c1-120"
done
for try in 1 2 3 4 5 6; do
  /usr/local/graft/bin/gpurun --timeout 2400 -- "$CMD" > /tmp/prof_all_$TAG.log 2>&1
  if grep -q "status=transient" /tmp/prof_all_$TAG.log; then sleep 100; continue; fi
  break
done
tail -8 /tmp/prof_all_$TAG.log
python tools/prof_extract.py $TAG
