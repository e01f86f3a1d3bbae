#!/bin/bash
# group API on however many GPUs the box has: tests, then bench through rt2_group_* at N = 1 and (if present) N = 2
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 1200 python -m pytest tests/test_gpu_group.py tests/test_gpu_speculate.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_group.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_group.log
tail -15 gpurun_out/pytest_group.log | cut -c1-300
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "rc=$?" >> gpurun_out/bench_n1.log
tail -3 gpurun_out/bench_n1.log | cut -c1-1500
N=$(nvidia-smi -L | wc -l)
if [ "$N" -ge 2 ]; then
  for n in 2 $([ "$N" -ge 4 ] && echo 4) $([ "$N" -ge 8 ] && echo 8); do
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench_n$n.log 2>&1; echo "rc=$?" >> gpurun_out/bench_n$n.log
    tail -2 gpurun_out/bench_n$n.log | cut -c1-900
  done
fi
