#!/bin/bash
# C5 frame time per library variant (gpurun_variants/*.so), both cameras
for v in "" $VARIANTS; do
  lib=""; [ -n "$v" ] && lib="gpurun_variants/$v.so"
  for cam in "" "--overhead"; do
    RT2_CUDA_LIB=$lib timeout 300 python tools/c5_frame.py --frames 6 $cam 2>&1 | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); ms=sorted(x['ms'] for x in d[2:]); r=d[-1]['rays']
    print('variant [$v] $cam', 'median ms', ms[len(ms)//2], 'Mrays/s', round(r/ms[len(ms)//2]/1e3,1), 'launches', d[-1]['launches'], 'schedule', d[-1]['schedule'])
except Exception as e: print('variant [$v] parse fail', e)"
  done
done
