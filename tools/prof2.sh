#!/bin/bash
# usage: tools/prof2.sh <tag>  -- ncu --set full capture of both decode-step kernels, per-region summaries in gpurun_out/
TAG=$1
/usr/local/graft/bin/gpurun --timeout 900 -- "for K in decode_scan_kernel decode_tail_kernel; do timeout 400 ncu --set full --clock-control none --import-source on -k regex:\$K -s 6 -c 1 -o gpurun_out/${TAG}_\$K python bench.py --no-cpu-baseline --steps 8 --warmup 3 > gpurun_out/ncu_${TAG}_\$K.log 2>&1; tail -1 gpurun_out/ncu_${TAG}_\$K.log | cut -c1-100; done" 2>&1 | tail -3
for K in decode_scan_kernel decode_tail_kernel; do
ncu -i gpurun_out/${TAG}_$K.ncu-rep --page raw --csv > gpurun_out/${TAG}_${K}_raw.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_$K.ncu-rep --page source --csv --print-source sass > gpurun_out/${TAG}_${K}_sass.csv 2>/dev/null
python tools/ncu_lines.py gpurun_out/${TAG}_${K}_sass.csv whisper_rs_b200/lib/libwrs_b200.so $K 2000 > gpurun_out/${TAG}_${K}_lines.txt
python tools/ncu_keys.py gpurun_out/${TAG}_${K}_raw.csv | grep -v "^----" > gpurun_out/${TAG}_${K}_keys.txt
grep "time_duration\|inst_executed.sum\|issue_active\|cycles_active\|registers" gpurun_out/${TAG}_${K}_keys.txt
done
