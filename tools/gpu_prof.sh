#!/bin/bash
# usage: gpu_prof.sh <name> <kernel regex> <prof_op args...>   -> gpurun_out/prof_<name>.ncu-rep
name=$1; rx=$2; shift 2
mkdir -p gpurun_out
python tools/prof_op.py "$@" > gpurun_out/plain_$name.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$rx -s 2 -c 1 -o gpurun_out/prof_$name -f python tools/prof_op.py "$@" > gpurun_out/ncu_$name.log 2>&1
tail -3 gpurun_out/ncu_$name.log
