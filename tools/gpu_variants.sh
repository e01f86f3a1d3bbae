#!/bin/bash
# parity on one trace variant, then a short bench per RT2_TRACE_VARIANT
mkdir -p gpurun_out
RT2_TRACE_VARIANT=${PARITY_VARIANT:-10} timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_golden.py -x -q 2>&1 | tail -3
for v in ${VARIANTS:-5 10 11 12 13 14 5}; do
  RT2_TRACE_VARIANT=$v timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('  v$v', round(d['value'],1), 'Mrays/s  trace ms/launch', round(d['roofline']['avg_launch_ms'],4), 'e2e', round(d['e2e']['value'],1))
except Exception as e: print('  v$v parse fail', e)"
done
