#!/bin/bash
# usage: gpu_prof1.sh <op> <kernel regex> [N] [prec]
op=$1; rx=$2; N=${3:-32}; prec=${4:-tf32}
mkdir -p gpurun_out
P="python tools/prof_op.py $op $N $prec"
$P > gpurun_out/plain_$op.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$rx -s 2 -c 1 -f -o gpurun_out/prof_$op $P > gpurun_out/ncu_$op.log 2>&1
tail -2 gpurun_out/ncu_$op.log
