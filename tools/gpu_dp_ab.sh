#!/bin/bash
# A/B of the SAH-optimal collapse (RT2_COLLAPSE_DP=1, default) against largest-child-first opening: parity subset, then bench with extras
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_adversarial_geometry.py tests/test_golden.py tests/test_gpu_anchors.py tests/test_gpu_speculate.py -m gpu -q -x > gpurun_out/pytest_dp.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_dp.log
tail -5 gpurun_out/pytest_dp.log
for dp in 1 0; do
  RT2_COLLAPSE_DP=$dp timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dp$dp.log 2>&1; echo "rc=$?" >> gpurun_out/bench_dp$dp.log
  python - <<PY
import json
l=[x for x in open('gpurun_out/bench_dp$dp.log') if x.startswith('{')]
if l:
    d=json.loads(l[-1]); ex=d.get('extra',{})
    print('DP=$dp C2', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), '| C5', ex.get('C5',{}).get('Mrays_s'), 'C5_overhead', ex.get('C5_overhead',{}).get('Mrays_s'), '| C3', ex.get('C3',{}).get('Mrays_s'), '| GI', ex.get('restir_gi_default_frame',{}).get('Mrays_s'), '| C1', ex.get('C1',{}).get('Mrays_s'))
else:
    print('DP=$dp: no line'); print(open('gpurun_out/bench_dp$dp.log').read()[-1500:])
PY
done
