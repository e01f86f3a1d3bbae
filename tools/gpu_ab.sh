#!/bin/bash
# A/B of alternative builds of librt2cuda (gpurun_variants/*.so, made by tools/build_variant.sh): short bench per build, default first and last
mkdir -p gpurun_out
run() { RT2_CUDA_LIB=$1 timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('  $2', round(d['value'],1), 'Mrays/s  trace ms/launch', round(d['roofline']['avg_launch_ms'],4), 'shade ms/launch', round(d['roofline']['avg_launch_ms']*(1-d['roofline']['trace_share_of_step'])/d['roofline']['trace_share_of_step'],4), 'e2e', round(d['e2e']['value'],1))
except Exception as e: print('  $2 parse fail', e)"; }
run "" default
for f in ${VARIANTS:-gpurun_variants/*.so}; do run $PWD/$f $(basename $f .so); done
run "" default
