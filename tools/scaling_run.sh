#!/bin/bash
# 1-8 GPU scaling on one box: bench.py (1080p weak, 4 spp per GPU) and the C4 probe (4K strong, 16 x 4 spp accumulated)
mkdir -p gpurun_out
port=29600
for n in 1 2 4 8; do
  port=$((port+1))
  if [ $n = 1 ]; then timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/scale_bench_n$n.json
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/scale_bench_n$n.json; fi
  python -c "
import json; d=json.load(open('gpurun_out/scale_bench_n$n.json')); print('bench N=$n', round(d['value'],1), 'Mrays/s e2e', round(d['e2e']['value'],1), 'ms/step', round(d['ms_per_step'],2))"
done
for n in 1 2 4 8; do
  port=$((port+1))
  if [ $n = 1 ]; then PYTHONPATH=. timeout 300 python tools/c4_probe.py 2>/dev/null | tail -1 > gpurun_out/scale_c4_n$n.json
  else PYTHONPATH=. timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port tools/c4_probe.py --check 2>/dev/null | tail -1 > gpurun_out/scale_c4_n$n.json; fi
  echo "c4 N=$n $(cut -c1-400 gpurun_out/scale_c4_n$n.json)"
done
