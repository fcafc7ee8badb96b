#!/bin/bash
# usage: tools/gpu_k2.sh <tag>  -- decode GPU tests + bench (decode section only); retries while the pod is busy
TAG=$1; shift
CMD="timeout 900 python -m pytest tests/test_gpu_decode.py tests/test_gpu_decode_loop.py tests/test_gpu_configs.py tests/test_decode_task.py -x -q -m gpu > gpurun_out/${TAG}_pytest.log 2>&1; echo \"pytest rc=\$?\"; tail -3 gpurun_out/${TAG}_pytest.log; timeout 600 python bench.py --no-cpu-baseline --only decode --steps 400 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo \"bench rc=\$?\"; tail -3 gpurun_out/${TAG}_bench.err"
for try in 1 2 3 4 5 6; do
  /usr/local/graft/bin/gpurun --timeout 1700 -- "$CMD" > /tmp/gpu_k2_$TAG.log 2>&1
  if grep -q "status=transient" /tmp/gpu_k2_$TAG.log; then sleep 100; continue; fi
  break
done
tail -9 /tmp/gpu_k2_$TAG.log
python - <<PY
import json
try:
    d=json.load(open('gpurun_out/${TAG}_bench.json'))
    for k in ('decode_step','decode_step_large','no_speech_probs','filter_apply'):
        print(k, 'ms', round(d[k]['ms_per_step'],5), 'frac', round(d[k]['roofline']['frac'],4), 'eager', d[k].get('eager_ms_per_step'))
except Exception as e: print('no bench', e)
PY
