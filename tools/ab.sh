#!/bin/bash
# on the GPU box: time every build/lib_*.so with kbench
for f in /root/repo/build/lib_*.so; do echo "== $(basename $f)"; ARENA_GPUACCT_LIB=$f python tools/kbench.py ${1:-20000000} | tail -3; done
