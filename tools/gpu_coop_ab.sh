#!/bin/bash
# eight-lanes-per-ray walk (RT2_COOP=1) against the lane-per-ray walk (RT2_COOP=0): C5 whole frame, one rank of 4 / 8, and the ReSTIR frames
for coop in 0 1; do
  for bw in 0 4 8; do
    RT2_COOP=$coop timeout 300 python tools/c5_frame.py --frames 6 --band-world $bw 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); ms=sorted(x['ms'] for x in d[2:]); m=ms[len(ms)//2]
print('coop=$coop C5 band_world $bw: median ms', m, 'Mrays/s', round(d[-1]['rays']/m/1e3,1))"
  done
  for what in c3 gi gi67; do
    RT2_COOP=$coop timeout 300 python tools/passes_frame.py $what --frames 8 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); ms=sorted(x['ms'] for x in d[2:]); m=ms[len(ms)//2]
print('coop=$coop $what median ms', m, 'Mrays/s', round(d[-1]['rays']/m/1e3,1))"
  done
done
