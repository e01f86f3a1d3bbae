#!/bin/bash
# A/B of alternative builds (gpurun_variants/*.so) on the ReSTIR frame variants (tools/gi_probe.py), ms per frame
export PYTHONPATH=.
run() { RT2_CUDA_LIB=$1 timeout 300 python tools/gi_probe.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('  $2', {k: v['ms_per_frame'] for k,v in d.items()})"; }
run "" default
for f in gpurun_variants/*.so; do run $PWD/$f $(basename $f .so); done
run "" default
