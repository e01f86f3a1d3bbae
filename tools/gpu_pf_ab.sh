#!/bin/bash
# flat-walk child prefetch A/B: C5 whole frame (both cameras) and one rank of 8, per library variant (gpurun_variants/*.so)
for v in "" pf1 pf2 pf3; do
  lib=""; [ -n "$v" ] && lib="gpurun_variants/$v.so"
  for cam in "" "--overhead"; do
    RT2_CUDA_LIB=$lib timeout 300 python tools/c5_frame.py --frames 6 $cam 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); ms=sorted(x['ms'] for x in d[2:]); print('variant [$v] $cam whole frame: median ms', ms[len(ms)//2])"
    RT2_CUDA_LIB=$lib timeout 300 python tools/c5_share_sweep.py --variants 1 $cam 2>&1 | tail -1 | cut -c1-120
  done
done
