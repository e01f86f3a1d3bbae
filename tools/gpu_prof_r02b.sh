#!/bin/bash
# r02b profiles (after the tree-quality work): launch list of the bench command + full capture of k_trace round-0 launches of the timed step
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra"
timeout 300 $CMD > gpurun_out/plain_c.log 2>&1 || exit 1
tail -1 gpurun_out/plain_c.log | cut -c1-200
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_trace|k_spec|k_path|k_primary|k_k1" -c 600 --csv --log-file gpurun_out/r02c_launches.csv $CMD > gpurun_out/ncu_list_c.log 2>&1; echo "list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_trace_dyn -s 80 -c 4 -o gpurun_out/r02c_prof_trace $CMD > gpurun_out/ncu_full_c.log 2>&1; echo "full rc=$?"
