#!/bin/bash
# scan_items_kernel time for every pods-per-warp setting (run under gpurun)
for L in 1 2 4 8 16 32; do
  ARENA_LOADER_LANES=$L ncu --metrics gpu__time_duration.sum --clock-control none -k regex:scan_items -c 3 --csv --log-file gpurun_out/lanes_$L.csv python tools/loader_timing.py --device --reps 1 --pods ${1:-10000} > /dev/null 2>&1
  echo "lanes=$L $(grep scan_items gpurun_out/lanes_$L.csv | tail -1 | awk -F'","' '{print $NF}')"
done
