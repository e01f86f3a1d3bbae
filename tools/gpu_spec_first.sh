#!/bin/bash
# first contact of speculative sample parallelism: parity, whole suite under RT2_SPECULATE=1, throughput per rank-of-N share
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_speculate.py -m gpu -q -x > gpurun_out/pytest_spec.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_spec.log
tail -30 gpurun_out/pytest_spec.log | cut -c1-300
RT2_SPECULATE=1 timeout 2400 python -m pytest tests -m gpu -q --deselect tests/test_gpu_speculate.py > gpurun_out/pytest_gpu_spec1.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_spec1.log
tail -12 gpurun_out/pytest_gpu_spec1.log | cut -c1-300
{
echo "== sequential (RT2_SPECULATE=0)"; RT2_SPECULATE=0 timeout 300 python tools/band_probe.py 0
echo "== speculative (RT2_SPECULATE=1)"; RT2_SPECULATE=1 timeout 300 python tools/band_probe.py 0
echo "== auto"; timeout 300 python tools/band_probe.py 0
} > gpurun_out/spec_sweep.txt 2>&1
cat gpurun_out/spec_sweep.txt
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench.log 2>&1; tail -1 gpurun_out/bench.log | cut -c1-700
