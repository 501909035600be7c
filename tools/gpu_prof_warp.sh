#!/bin/bash
mkdir -p gpurun_out
P="python tools/prof_op.py warp_bwd 32 tf32"
$P > gpurun_out/plain_warp_bwd.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:warp_tile_scatter -s 2 -c 1 -f -o gpurun_out/prof_warp_scatter $P > gpurun_out/ncu_warp_scatter.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:warp_tile_gather -s 2 -c 1 -f -o gpurun_out/prof_warp_gflow $P > gpurun_out/ncu_warp_gflow.log 2>&1
P="python tools/prof_op.py warp_fwd 32 tf32"
$P > gpurun_out/plain_warp_fwd.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:warp_tile_gather -s 2 -c 1 -f -o gpurun_out/prof_warp_fwd $P > gpurun_out/ncu_warp_fwd.log 2>&1
tail -1 gpurun_out/ncu_warp_*.log
