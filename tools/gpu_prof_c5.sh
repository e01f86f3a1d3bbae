#!/bin/bash
# ncu --set full of the C5 (flattened instancing) trace kernel: bounce launches of the second frame
mkdir -p gpurun_out
CMD="python tools/c5_frame.py --frames 2"
timeout 300 $CMD > gpurun_out/c5_plain.log 2>&1 || { tail -5 gpurun_out/c5_plain.log; exit 1; }
tail -1 gpurun_out/c5_plain.log | cut -c1-600
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_c5_launches.csv $CMD > gpurun_out/c5_ncu_list.log 2>&1; echo "list rc=$?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_trace -s ${SKIP:-20} -c 6 -o gpurun_out/r02_prof_c5_trace $CMD > gpurun_out/c5_ncu_full.log 2>&1; echo "full rc=$?"
ls -la gpurun_out/*.ncu-rep | tail -3
