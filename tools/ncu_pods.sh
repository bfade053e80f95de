#!/bin/bash
# usage (on the GPU box): tools/ncu_pods.sh <tag> [pods]
# plain run first (must exit 0), then one full ncu capture of the pods kernel.
TAG=${1:-x}; PODS=${2:-20000000}
python tools/kbench.py $PODS $((PODS/100)) 3 > gpurun_out/kbench_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:top_node_pods_kernel -s 9 -c 1 -f -o gpurun_out/prof_$TAG \
    python tools/kbench.py $PODS $((PODS/100)) 3 > gpurun_out/ncu_$TAG.log 2>&1
tail -2 gpurun_out/kbench_$TAG.log
