#!/bin/bash
# usage (on the GPU box): tools/debug/ab.sh  -- bench every tools/debug/var_*.so variant of the library, print the decode-step time
cp whisper_rs_b200/lib/libwrs_b200.so /tmp/orig.so
for v in tools/debug/var_*.so; do
  cp $v whisper_rs_b200/lib/libwrs_b200.so
  python bench.py --no-cpu-baseline --steps 400 --warmup 5 > /tmp/o.json 2>/tmp/e.txt || tail -3 /tmp/e.txt
  python - "$v" <<'P'
import json,sys
d=json.loads(open('/tmp/o.json').read().strip().splitlines()[-1])
print(sys.argv[1], 'logmel ms', round(d['ms_per_step'],4), 'decode us', round(d['decode_step']['ms_per_step']*1e3,2), 'nospeech us', round(d['no_speech_probs']['ms_per_step']*1e3,2))
P
done
cp /tmp/orig.so whisper_rs_b200/lib/libwrs_b200.so
