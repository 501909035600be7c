#!/bin/bash
# launch list of one bench step + full ncu capture of the two tensor-core correlation kernels
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --precision tf32 --no-cpu-baseline --no-ref-gpu --no-e2e"
$B > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 80 -c 40 --csv --log-file gpurun_out/launches_tf32.csv $B > gpurun_out/ncu_launches.log 2>&1
P="python tools/prof_corr.py tf32 bwd 32"
$P > gpurun_out/plain_bwd.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:corr_tc_bwd -s 2 -c 1 -o gpurun_out/prof_bwd -f $P > gpurun_out/ncu_bwd.log 2>&1
P="python tools/prof_corr.py tf32 fwd 32"
$P > gpurun_out/plain_fwd.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:corr_tc_fwd -s 2 -c 1 -o gpurun_out/prof_fwd -f $P > gpurun_out/ncu_fwd.log 2>&1
ls -la gpurun_out | tail; tail -3 gpurun_out/ncu_bwd.log
