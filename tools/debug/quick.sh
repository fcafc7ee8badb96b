#!/bin/bash
# usage (on the GPU box): tools/debug/quick.sh  -- GPU parity tests + a short bench, one summary line
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --no-cpu-baseline --steps 400 --warmup 5 2>/tmp/bench.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('logmel us', round(d['ms_per_step']*1e3,2), 'decode us', round(d['decode_step']['ms_per_step']*1e3,2), 'nospeech us', round(d['no_speech_probs']['ms_per_step']*1e3,2), 'kv ms', round(d['kv_reorder']['ms_per_step'],3), 'e2e seg/s', round(d['e2e']['value']))" || tail -5 /tmp/bench.err
