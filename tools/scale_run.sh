#!/bin/bash
# weak scaling at 2/4/8 GPUs + strong at 8 (run under gpurun --gpus 8); writes gpurun_out/scale_*.json
PORT=29700
for cfg in "2 weak" "4 weak" "8 weak" "8 strong"; do
  set -- $cfg; N=$1; S=$2; PORT=$((PORT+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $N --steps 100 --warmup 5 --no-cpu --scaling $S > gpurun_out/scale_${S}_g$N.json 2> gpurun_out/scale_${S}_g$N.err
  python - <<PY || tail -5 gpurun_out/scale_${S}_g$N.err
import json
d = json.load(open("gpurun_out/scale_${S}_g$N.json"))
print("$S N=$N", "%.4g pods/s" % d["value"], "%.4f ms/step" % d["ms_per_step"], "kernel %.4f ms" % d["roofline"]["kernel_ms"],
      "e2e %.4g (%.2f ms)" % (d["e2e"]["value"], d["e2e"]["ms_per_step"]), "wide %.3g" % d["e2e"]["wide"]["value"])
PY
done
