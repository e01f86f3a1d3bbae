#!/bin/bash
# C3 / GI / GI+K6K7 frame times per library variant
for v in "" $VARIANTS; do
  lib=""; [ -n "$v" ] && lib="gpurun_variants/$v.so"
  for what in c3 gi gi67; do
    RT2_CUDA_LIB=$lib timeout 300 python tools/passes_frame.py $what --frames 8 2>&1 | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); ms=sorted(x['ms'] for x in d[2:]); m=ms[len(ms)//2]
    print('variant [$v] $what median ms', m, 'Mrays/s', round(d[-1]['rays']/m/1e3,1))
except Exception as e: print('variant [$v] $what parse fail', e)"
  done
done
