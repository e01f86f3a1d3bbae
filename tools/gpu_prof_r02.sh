#!/bin/bash
# r02 profiles: full captures of k_trace (round-0 launches of the timed step: 32 sequential warm-up launches + 2 x 24 speculative
# ones come first) and k_spec_shade (2 x 24 come first)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 || exit 1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_trace_dyn -s 80 -c 4 -o gpurun_out/r02_prof_trace $CMD > gpurun_out/ncu_full.log 2>&1; echo "full rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_spec_shade -s 48 -c 4 -o gpurun_out/r02_prof_shade $CMD > gpurun_out/ncu_full2.log 2>&1; echo "full2 rc=$?"
