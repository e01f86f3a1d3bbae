#!/bin/bash
# ncu --set full of every kernel of one C3 frame (K3 + K9) after the visibility changes (reuse of the initial ray, any-hit batches)
mkdir -p gpurun_out
CMD="python tools/passes_frame.py c3 --frames 3"
timeout 300 $CMD > gpurun_out/passes_c3_d.log 2>&1 || { tail -3 gpurun_out/passes_c3_d.log; exit 1; }
tail -1 gpurun_out/passes_c3_d.log | cut -c1-300
n=$(python -c "import json; print(json.loads(open('gpurun_out/passes_c3_d.log').read().strip().split('\n')[-1])[-1]['launches'])")
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_gbuffer|k_di_|k_trace|k_brdf|k_shade_sec|k_gi_" -s $((2 * n)) -c $n -o gpurun_out/r02d_prof_c3 $CMD > gpurun_out/ncu_c3_d.log 2>&1; echo "c3: $n launches per frame, ncu rc=$?"
ncu -i gpurun_out/r02d_prof_c3.ncu-rep --page raw --csv > gpurun_out/r02d_prof_c3_raw.csv 2>/dev/null && rm -f gpurun_out/r02d_prof_c3.ncu-rep
ls -la gpurun_out/r02d_prof_c3*
