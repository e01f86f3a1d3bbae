#!/bin/bash
# GPU contact: smoke, gpu tests, bench, then the ncu launch list + one full capture of the top kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
tail -5 gpurun_out/smoke.log; tail -30 gpurun_out/pytest_gpu.log; tail -5 gpurun_out/bench.log
if [ "$1" = "ncu" ]; then
  CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
  timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
  echo "ncu list rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_trace -s 20 -c 3 -o gpurun_out/prof_trace $CMD > gpurun_out/ncu_full.log 2>&1
  echo "ncu full rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_k1_shade -s 20 -c 2 -o gpurun_out/prof_shade $CMD > gpurun_out/ncu_full2.log 2>&1
  echo "ncu full2 rc=$?"
fi
