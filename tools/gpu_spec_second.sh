#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_speculate.py tests/test_gpu_parity.py tests/test_golden.py -m gpu -q -x > gpurun_out/pytest_spec.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_spec.log
tail -12 gpurun_out/pytest_spec.log | cut -c1-300
RT2_SPECULATE=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_golden.py tests/test_gpu_anchors.py -m gpu -q -x > gpurun_out/pytest_spec1.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_spec1.log
tail -5 gpurun_out/pytest_spec1.log | cut -c1-300
{
echo "== sequential (RT2_SPECULATE=0)"; RT2_SPECULATE=0 timeout 300 python tools/band_probe.py 0
echo "== speculative (RT2_SPECULATE=1)"; RT2_SPECULATE=1 timeout 300 python tools/band_probe.py 0
echo "== speculative, chains 1,2,3,4"; RT2_SPECULATE=1 timeout 600 python tools/band_probe.py 1,2,3,4
} > gpurun_out/spec_sweep2.txt 2>&1
cat gpurun_out/spec_sweep2.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench.log 2>&1; tail -1 gpurun_out/bench.log | cut -c1-330
RT2_SPECULATE=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/spec_launches_n8.csv python tools/spec_profile.py 8 1 > gpurun_out/specprof_ncu_8.log 2>&1
echo "== N=8 speculative, last frame"; python tools/ncu_kernel_totals.py gpurun_out/spec_launches_n8.csv 31
