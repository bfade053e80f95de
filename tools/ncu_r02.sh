#!/bin/bash
# Round-2 ncu evidence (on the GPU box, one GPU): launch list of the bench command, one full capture each of the
# top-node kernel, the packed kernel and the bulk-copy staged variant.  Every command runs plain first.
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-wire"
$B > gpurun_out/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches.csv $B > gpurun_out/ncu_list.log 2>&1
$B > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:top_node_pods_kernel -s 3 -c 2 -f -o gpurun_out/r02_pods $B > gpurun_out/ncu_pods.log 2>&1
$B > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:top_node_packed_kernel -s 12 -c 1 -f -o gpurun_out/r02_packed $B > gpurun_out/ncu_packed.log 2>&1
S="python tools/kbench.py 20000000 200000 3"
ARENA_STAGED=1 KBENCH_NOPARITY=1 $S > gpurun_out/ncu_plain2.log 2>&1 &&
ARENA_STAGED=1 KBENCH_NOPARITY=1 ncu --set full --clock-control none --import-source on -k regex:staged -s 5 -c 1 -f -o gpurun_out/r02_staged $S > gpurun_out/ncu_staged.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/r02_launches.csv
