#!/bin/bash
# ncu evidence for the path tracer: launch list of the bench command + one full capture of three k_trace launches and two k_k1_shade launches
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_trace -s 20 -c 3 -o gpurun_out/prof_trace $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_k1_shade -s 20 -c 2 -o gpurun_out/prof_shade $CMD > gpurun_out/ncu_shade.log 2>&1
echo "ncu shade rc=$?"
ls -la gpurun_out/*.ncu-rep
