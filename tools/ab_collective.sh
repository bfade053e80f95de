#!/bin/bash
# A/B of the multi-GPU totals exchange: NCCL all-reduce vs in-kernel peer exchange (run under gpurun --gpus N)
N=${1:-2}; shift
PORT=29600
for c in nccl fused posted; do
  PORT=$((PORT+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $N --steps 200 --warmup 5 --no-cpu --e2e-steps 2 --collective $c "$@" > gpurun_out/ab_${c}_g$N.json 2> gpurun_out/ab_${c}_g$N.err
  python - <<PY || tail -5 gpurun_out/ab_${c}_g$N.err
import json
d = json.load(open("gpurun_out/ab_${c}_g$N.json"))
print("$c", "N=$N", "%.4g pods/s" % d["value"], "%.4f ms/step" % d["ms_per_step"], "kernel %.4f ms" % d["roofline"]["kernel_ms"], "|", d["collective"])
PY
done
