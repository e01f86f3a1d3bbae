#!/bin/bash
# round 2, after the primary-hit reuse, on an 8-GPU box: the bench step at N = 8 (weak) and strong scaling of C2-strong / C5 at N = 1, 8
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l); echo "GPUs: $N"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29733 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline 2>gpurun_out/weak_d_n$N.err | tail -1 > gpurun_out/weak_d_n$N.json
python -c "
import json
d=json.load(open('gpurun_out/weak_d_n$N.json')); print('weak N=$N', round(d['value'],1), 'Mrays/s  e2e', round(d['e2e']['value'],1), ' ms/step', round(d['ms_per_step'],3), ' clocks', d['clocks']['sm_mhz'], d['clocks']['reasons'])"
timeout 900 python tools/strong_scaling.py --ns 1,$N --workloads c2s,c5 > gpurun_out/strong_d_n$N.log 2>&1; echo "strong rc=$?"
grep -v "^{" gpurun_out/strong_d_n$N.log | grep -v NCCL | cut -c1-260
