#!/bin/bash
# A/B of variant builds: tools/gpu_ab.sh "<time_ops args>" lib1.so lib2.so ...
args=$1; shift
echo "== default"; timeout 300 python tools/time_ops.py $args
for l in "$@"; do echo "== $l"; FN2_OPS_LIBRARY=$PWD/$l timeout 300 python tools/time_ops.py $args; done
