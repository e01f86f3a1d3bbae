#!/bin/bash
# run the given pytest selection on the GPU box: tools/gpu_quick.sh <pytest args>
mkdir -p gpurun_out
timeout 1500 python -m pytest "$@" -m gpu -q -x > gpurun_out/pytest_quick.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_quick.log
tail -30 gpurun_out/pytest_quick.log | cut -c1-400
