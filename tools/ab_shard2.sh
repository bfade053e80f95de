#!/bin/bash
# 2 GPUs, 2.5e7 pods in total = the 8-GPU share (1.25e7) per rank, strong split, posted exchange: library A vs B
for lib in "$@"; do
  for rep in 1 2; do
    ARENA_GPUACCT_LIB=$lib timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29800 + RANDOM % 100)) \
      bench.py --gpus 2 --steps 300 --warmup 10 --no-cpu --no-wire --pods 25000000 --nodes 250000 --e2e-steps 2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$lib', 'ms/step %.4f' % d['ms_per_step'], 'kernel %.4f' % d['roofline']['kernel_ms'], 'parity', d['parity_ok'])"
  done
done
