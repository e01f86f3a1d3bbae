#!/bin/bash
# A/B of trace-kernel variants: parity tests on the default, then a short bench per variant
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
for v in ${VARIANTS:-0 1 2 3 4 5 6 7}; do
  RT2_TRACE_VARIANT=$v timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v$v.log 2>&1
  echo "variant $v rc=$?"; python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/bench_v$v.log") if x.startswith("{")][-1]; d=json.loads(l)
    print("  v$v", round(d["value"],1), "Mrays/s  trace_ms/launch", round(d["roofline"]["avg_launch_ms"],4), "e2e", round(d["e2e"]["value"],1))
except Exception as e: print("  parse fail", e)
PY
done
