#!/bin/bash
# Round checks (smoke, gpu tests, bench) then the ncu evidence: launch list of the bench command, one full capture of the
# trace / shade kernels, and one of the ReSTIR wavefront kernels.  Every ncu command runs only after the same command exited 0 alone.
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench.log | cut -c1-400
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_trace -s 20 -c 3 -o gpurun_out/prof_trace $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
CMD2="python tools/gi_probe.py"
export PYTHONPATH=.
timeout 300 $CMD2 > gpurun_out/gi_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_di_stage|k_brdf_rays|k_shade_secondary|k_gi_|k_trace|k_gbuffer' -s 40 -c 16 -o gpurun_out/prof_restir $CMD2 > gpurun_out/ncu_full3.log 2>&1
echo "ncu restir rc=$?"
ls -la gpurun_out/*.ncu-rep
