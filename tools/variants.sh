#!/bin/bash
# build A/B variants of the library: tools/variants.sh name "-DFLAG=.. -DFLAG2=.." [name2 "flags2" ...]
cd /root/repo
while [ $# -gt 1 ]; do
  name=$1; flags=$2; shift 2
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC $flags -Xptxas -v -c arena_b200/csrc/gpuacct.cu -o /tmp/gpuacct_$name.o 2>&1 | grep -A2 "top_node_pods" | grep -E "Used" | sed "s/^/$name: /"
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/lib_$name.so /tmp/gpuacct_$name.o $(ls arena_b200/_obj/*.o | grep -v gpuacct.o) -lcudart
done
