#!/bin/bash
# A/B of the BVH2 reinsertion rounds (RT2_REINSERT_ITERS): parity subset with the default, then build time + C5 frame + bench per setting
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_adversarial_geometry.py tests/test_golden.py tests/test_gpu_anchors.py tests/test_gpu_speculate.py -m gpu -q -x > gpurun_out/pytest_ri.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_ri.log
tail -4 gpurun_out/pytest_ri.log
for it in ${ITERS:-0 12}; do
  echo "== RT2_REINSERT_ITERS=$it"
  RT2_REINSERT_ITERS=$it timeout 300 python tools/c5_frame.py --frames 6 2>&1 | tail -2 | python -c "
import json,sys
l=sys.stdin.read().strip().split('\n'); print(l[0])
d=json.loads(l[-1]); ms=sorted(x['ms'] for x in d[2:]); print('  C5 median ms', ms[len(ms)//2], 'Mrays/s', round(d[-1]['rays']/ms[len(ms)//2]/1e3,1))"
  RT2_REINSERT_ITERS=$it timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_ri$it.log 2>&1
  python - <<PY
import json
l=[x for x in open('gpurun_out/bench_ri$it.log') if x.startswith('{')]
if l:
    d=json.loads(l[-1]); ex=d.get('extra',{})
    print('  bench C2', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), '| C5', ex.get('C5',{}).get('Mrays_s'), 'C5_overhead', ex.get('C5_overhead',{}).get('Mrays_s'), '| C3', ex.get('C3',{}).get('Mrays_s'), '| GI', ex.get('restir_gi_default_frame',{}).get('Mrays_s'))
else:
    print('  no line'); print(open('gpurun_out/bench_ri$it.log').read()[-1500:])
PY
done
