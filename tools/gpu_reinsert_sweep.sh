#!/bin/bash
for it in ${ITERS:-12 24 48}; do
  echo "== RT2_REINSERT_ITERS=$it"
  for cam in "" "--overhead"; do
  RT2_REINSERT_ITERS=$it timeout 300 python tools/c5_frame.py --frames 6 $cam 2>&1 | tail -2 | python -c "
import json,sys
l=sys.stdin.read().strip().split('\n'); print(' ', l[0])
d=json.loads(l[-1]); ms=sorted(x['ms'] for x in d[2:]); print('  C5 $cam median ms', ms[len(ms)//2], 'Mrays/s', round(d[-1]['rays']/ms[len(ms)//2]/1e3,1))"
  done
  RT2_REINSERT_ITERS=$it timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extra 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('  bench C2', round(d['value'],1), 'e2e', round(d['e2e']['value'],1))"
done
