#!/bin/bash
# round-2 first GPU contact: full gpu test suite (incl. full-size parity + anchors), sanitizers, baseline bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1; nproc >> gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -q -x --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
for tool in memcheck racecheck initcheck; do
  RT2_CHAINS=8 timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize.py > gpurun_out/sanitizer_$tool.txt 2>&1
  echo "$tool rc=$?" >> gpurun_out/sanitizer_$tool.txt
  tail -4 gpurun_out/sanitizer_$tool.txt
done
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
tail -3 gpurun_out/bench.log
