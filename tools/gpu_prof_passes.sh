#!/bin/bash
# ncu --set full of the ReSTIR pass kernels: one whole frame (the third) of C3 (K3 + K9), of the default GI frame and of GI with K6 + K7
mkdir -p gpurun_out
for what in c3 gi gi67; do
  CMD="python tools/passes_frame.py $what --frames 3"
  timeout 300 $CMD > gpurun_out/passes_$what.log 2>&1 || { tail -3 gpurun_out/passes_$what.log; continue; }
  tail -1 gpurun_out/passes_$what.log | cut -c1-300
  n=$(python -c "import json; print(json.loads(open('gpurun_out/passes_$what.log').read().strip().split('\n')[-1])[-1]['launches'])")
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_gbuffer|k_di_|k_trace|k_brdf|k_shade_sec|k_gi_" -s $((2 * n)) -c $n -o gpurun_out/r02b_prof_$what $CMD > gpurun_out/ncu_$what.log 2>&1; echo "$what: $n launches per frame, ncu rc=$?"
  # the reports are 25-33 MB each (gpurun brings back at most 64 MB): keep the raw page as CSV, drop the report
  ncu -i gpurun_out/r02b_prof_$what.ncu-rep --page raw --csv > gpurun_out/r02b_prof_${what}_raw.csv 2>/dev/null && rm -f gpurun_out/r02b_prof_$what.ncu-rep
done
ls -la gpurun_out/r02b_prof_*
