#!/bin/bash
# usage (on the GPU box, library built with WRS_NVCC_EXTRA=-DWRS_TRACE): tools/debug/split.sh -- decode step time: both launches, scan alone, tail alone (CUDA graph, bench.py)
for m in both scan tail; do
  if [ $m = both ]; then unset WRS_DEBUG_ONLY; else export WRS_DEBUG_ONLY=$m; fi
  python bench.py --no-cpu-baseline --steps 400 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$m', 'decode us', round(d['decode_step']['ms_per_step']*1e3,2), 'eager us', round(d['decode_step']['eager_ms_per_step']*1e3,2))"
done
