#!/bin/bash
# round 2, on an N-GPU box: rt2_group tests at every device count, bench.py weak scaling through torchrun, strong scaling in one process
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l); echo "GPUs: $N"
timeout 1500 python -m pytest tests/test_gpu_group.py -m gpu -q > gpurun_out/pytest_group_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_group_n$N.log
tail -4 gpurun_out/pytest_group_n$N.log | cut -c1-300
port=29700
for n in 1 2 4 8; do
  [ $n -le $N ] || continue
  port=$((port+1))
  if [ $n = 1 ]; then timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extra 2>gpurun_out/weak_n$n.err | tail -1 > gpurun_out/weak_n$n.json
  else timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps 10 --warmup 3 --no-cpu-baseline 2>gpurun_out/weak_n$n.err | tail -1 > gpurun_out/weak_n$n.json; fi
  python -c "
import json
try:
    d=json.load(open('gpurun_out/weak_n$n.json')); print('weak N=$n', round(d['value'],1), 'Mrays/s  e2e', round(d['e2e']['value'],1), ' ms/step', round(d['ms_per_step'],3), ' clocks', d['clocks']['sm_mhz'], d['clocks']['reasons'])
except Exception as e: print('weak N=$n failed', e)"
done
timeout 1200 python tools/strong_scaling.py --ns 1,2,4,8 > gpurun_out/strong_n$N.log 2>&1; echo "strong rc=$?"
grep -v "^{" gpurun_out/strong_n$N.log | grep -v NCCL | cut -c1-260
