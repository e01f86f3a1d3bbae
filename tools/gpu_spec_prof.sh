#!/bin/bash
mkdir -p gpurun_out
for n in 1 8; do
  RT2_SPECULATE=1 timeout 300 python tools/spec_profile.py $n 2 > gpurun_out/specprof_plain_$n.log 2>&1
  RT2_SPECULATE=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/spec_launches_n$n.csv python tools/spec_profile.py $n 1 > gpurun_out/specprof_ncu_$n.log 2>&1
  echo "== N=$n speculative, last frame (32 launches)"; python tools/ncu_kernel_totals.py gpurun_out/spec_launches_n$n.csv 32
done
RT2_SPECULATE=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/seq_launches_n8.csv python tools/spec_profile.py 8 0 > gpurun_out/seqprof_ncu_8.log 2>&1
echo "== N=8 sequential, one frame"; python tools/ncu_kernel_totals.py gpurun_out/seq_launches_n8.csv
