#!/bin/bash
# BASELINE configs[3] (one 1e8-pod snapshot cut by node range) at the rank counts given (default 2 4 8), then configs[4]
# (skewed) and the weak-scaling and `top job` arms at the largest.  Run under gpurun --gpus N; writes gpurun_out/scale_*.json
NS=${*:-2 4 8}
PORT=29700
run() {   # name N args...
  name=$1; N=$2; shift 2; PORT=$((PORT+1))
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $N --steps 200 --warmup 5 --no-cpu --e2e-steps 4 --graph "$@" > gpurun_out/scale_${name}_g$N.json 2> gpurun_out/scale_${name}_g$N.err
  python - <<PY || tail -5 gpurun_out/scale_${name}_g$N.err
import json
d = json.load(open("gpurun_out/scale_${name}_g$N.json"))
r, e = d.get("roofline"), d.get("e2e")
print("$name N=$N", "%.4g pods/s" % d["value"], "%.4f ms/step" % d["ms_per_step"], "parity_ok", d.get("parity_ok"),
      ("kernel %.4f ms frac %.3f" % (r["kernel_ms"], r["frac"])) if r else "",
      ("e2e %.4g (%.2f ms) wide %.3g" % (e["value"], e["ms_per_step"], e["wide"]["value"])) if e else "",
      "imbalance %.4f" % d["config"].get("shard_imbalance_max_over_mean", 1.0), "graph", d.get("graph_ms_per_step"))
PY
}
LAST=1
for N in $NS; do run strong $N; LAST=$N; done
# SCALE_EXTRA: the other arms at the largest rank count (default all three; "" = strong only)
for arm in ${SCALE_EXTRA-skewed weak job}; do
  case $arm in
    skewed) run skewed $LAST --skewed ;;
    weak) run weak $LAST --scaling weak ;;
    job) run job $LAST --command job ;;
  esac
done
