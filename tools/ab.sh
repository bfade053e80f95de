#!/bin/bash
# on the GPU box: parity + timing of every build/lib_*.so with kbench (pods kernel) -- tools/variants.sh builds them
for f in /root/repo/build/lib_*.so; do echo "== $(basename $f)"; ARENA_GPUACCT_LIB=$f python tools/kbench.py ${1:-100000000} ${2:-1000000} 50 2>&1 | grep -E "MISMATCH|pods_kernel|Error|error" | tail -3; done
