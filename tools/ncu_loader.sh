#!/bin/bash
# One `ncu --set full` capture of the GPU loader's per-pod kernel (run under gpurun, one GPU).
#   tools/ncu_loader.sh [pods]   -> gpurun_out/loader_scan_items.ncu-rep
set -e
PODS=${1:-10000}
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:scan_items_kernel -s 2 -c 1 \
    -o gpurun_out/loader_scan_items -f env ARENA_LOADER_LANES=${LANES:-0} python tools/loader_timing.py --device --pods $PODS --reps 1 > gpurun_out/ncu_loader.log 2>&1
tail -2 gpurun_out/ncu_loader.log | cut -c1-300
