#!/bin/bash
# usage: tools/gpu_k1.sh <tag> [extra pytest files]  -- log-mel GPU tests, bench line, ncu --set full of the log-mel kernel; retries while the pod is busy
TAG=$1; shift
CMD="timeout 600 python -m pytest tests/test_gpu_logmel.py tests/test_gpu_configs.py tests/test_zz_no_alloc.py $* -x -q -m gpu > gpurun_out/${TAG}_pytest.log 2>&1; echo \"pytest rc=\$?\"; tail -4 gpurun_out/${TAG}_pytest.log; timeout 600 python bench.py --no-cpu-baseline --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err && ncu --set full --clock-control none --import-source on -k regex:logmel_kernel -s 4 -c 1 -o gpurun_out/${TAG}_logmel python bench.py --no-cpu-baseline --steps 8 --warmup 3 > gpurun_out/${TAG}_ncu.log 2>&1; tail -1 gpurun_out/${TAG}_ncu.log | cut -c1-200"
for try in 1 2 3 4 5 6; do
  /usr/local/graft/bin/gpurun --timeout 1200 -- "$CMD" > /tmp/gpu_k1_$TAG.log 2>&1
  if grep -q "status=transient" /tmp/gpu_k1_$TAG.log; then sleep 120; continue; fi
  break
done
tail -7 /tmp/gpu_k1_$TAG.log
python - <<PY
import json
try:
    d=json.load(open('gpurun_out/${TAG}_bench.json'))
    print('K1 ms_per_step', d['ms_per_step'], 'frac', d['roofline']['frac'], 'e2e', d['e2e']['value'], 'decode ms', d['decode_step']['ms_per_step'])
except Exception as e: print('no bench', e)
PY
ncu -i gpurun_out/${TAG}_logmel.ncu-rep --page raw --csv > gpurun_out/${TAG}_logmel_raw.csv 2>/dev/null
python tools/ncu_keys.py gpurun_out/${TAG}_logmel_raw.csv | sed -n 5,40p
ncu -i gpurun_out/${TAG}_logmel.ncu-rep --page source --csv --print-source sass > gpurun_out/${TAG}_logmel_sass.csv 2>/dev/null
