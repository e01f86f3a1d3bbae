#!/bin/bash
# r02d profiles (after primary-hit reuse): launch list of the bench command + full capture of the first four k_trace launches of the timed
# step (speculative schedule, 2 chains: the representatives' primary walk of each chain, then bounce 1 of each chain)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra"
timeout 300 $CMD > gpurun_out/plain_d.log 2>&1 || exit 1
tail -1 gpurun_out/plain_d.log | cut -c1-200
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_trace|k_spec|k_path|k_primary|k_k1" -c 400 --csv --log-file gpurun_out/r02d_launches.csv $CMD > gpurun_out/ncu_list_d.log 2>&1; echo "list rc=$?"
# warm-up: frame 1 sequential (2 chains x 13 walks), frames 2-3 speculative (2 chains x 12 walks each) , the fix-up and candidate rounds included = 74 k_trace launches before the timed step
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_trace_dyn -s 74 -c 4 -o gpurun_out/r02d_prof_trace $CMD > gpurun_out/ncu_full_d.log 2>&1; echo "full rc=$?"
ncu -i gpurun_out/r02d_prof_trace.ncu-rep --page raw --csv > gpurun_out/r02d_k_trace_full_raw.csv 2>/dev/null
ls -la gpurun_out/r02d_*
[ $(stat -c %s gpurun_out/r02d_prof_trace.ncu-rep) -gt 40000000 ] && rm -f gpurun_out/r02d_prof_trace.ncu-rep
