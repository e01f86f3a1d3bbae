#!/bin/bash
# C5 (instanced scene, K1 + K13, RGBA8 band gather) on 1/2/4/8 GPUs of one box (NS overrides the list)
mkdir -p gpurun_out; port=29800
for n in ${NS:-1 2 4 8}; do port=$((port+1))
  if [ $n = 1 ]; then PYTHONPATH=. timeout 300 python tools/c5_probe.py 2>/dev/null | tail -1 > gpurun_out/scale_c5_n$n.json
  else PYTHONPATH=. timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port tools/c5_probe.py --check 2>/dev/null | tail -1 > gpurun_out/scale_c5_n$n.json; fi
  python -c "
import json; d=json.load(open('gpurun_out/scale_c5_n$n.json')); print('C5 N=$n', d['ms_per_frame'], 'ms/frame', d['Mrays_s'], 'Mrays/s', d.get('sharded_equals_whole_frame_bit_exact'))"
done
