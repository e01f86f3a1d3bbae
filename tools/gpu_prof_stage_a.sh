#!/bin/bash
# source-level ncu capture of K9's stage A (local-light RIS) on C3
mkdir -p gpurun_out
CMD="python tools/passes_frame.py c3 --frames 3"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_di_stage_a" -s 2 -c 1 -o gpurun_out/r02d_prof_stage_a $CMD > gpurun_out/ncu_stage_a.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r02d_prof_stage_a.ncu-rep --page source --csv --print-source sass > gpurun_out/r02d_stage_a_source_sass.csv 2>/dev/null
ncu -i gpurun_out/r02d_prof_stage_a.ncu-rep --page source --csv --print-source cuda > gpurun_out/r02d_stage_a_source_cuda.csv 2>/dev/null
ncu -i gpurun_out/r02d_prof_stage_a.ncu-rep --page raw --csv > gpurun_out/r02d_stage_a_raw.csv 2>/dev/null
ls -la gpurun_out/r02d_*stage_a*
