#!/bin/bash
# the no-alloc entry point's test: alone in a fresh process, then at the end of the whole gpu suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/pytest_noalloc_alone.log
tail -25 gpurun_out/pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -q -x --durations=5 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -14 gpurun_out/pytest_gpu.log
