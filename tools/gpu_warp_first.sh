#!/bin/bash
# first contact of the warp-local wavefront kernel (RT2_PATH_MODE=1): parity, then throughput sweeps at N=1 and rank-of-N sizes
mkdir -p gpurun_out
timeout 200 python tools/c5_bf_debug.py > gpurun_out/c5_bf_debug.txt 2>&1; tail -12 gpurun_out/c5_bf_debug.txt
RT2_PATH_MODE=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_golden.py tests/test_gpu_anchors.py "tests/test_gpu_fullsize.py::test_c2_full_size_whole_frame" -m gpu -q -x > gpurun_out/pytest_mode1.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_mode1.log
tail -8 gpurun_out/pytest_mode1.log
{
echo "== band_probe mode 0 (launch per bounce)"; timeout 300 python tools/band_probe.py 0
for P in 0 64 128 256 512; do for B in 0 5 4; do
  echo "== band_probe mode 1 P=$P blocks=$B"; RT2_PATH_MODE=1 RT2_WARP_P=$P RT2_WARP_BLOCKS=$B timeout 300 python tools/band_probe.py 0
done; done
} > gpurun_out/warp_sweep.txt 2>&1
cat gpurun_out/warp_sweep.txt
