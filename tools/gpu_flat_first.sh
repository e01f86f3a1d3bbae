#!/bin/bash
# first contact of flattened instancing: full gpu suite (flat on by default), two-level subset, C5 throughput in each mode
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -16 gpurun_out/pytest_gpu.log
RT2_USE_FLAT=0 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -q -x -k "instanced or c5" > gpurun_out/pytest_twolevel.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_twolevel.log
tail -4 gpurun_out/pytest_twolevel.log
{
for mode in 0 1; do for flat in 0 1; do
  echo "== C5 RT2_PATH_MODE=$mode RT2_USE_FLAT=$flat"; RT2_PATH_MODE=$mode RT2_USE_FLAT=$flat timeout 300 python tools/c5_probe.py --frames 10 | tail -1
done; done
} > gpurun_out/c5_flat.txt 2>&1
cat gpurun_out/c5_flat.txt
