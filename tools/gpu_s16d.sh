#!/bin/bash
# bench line + ncu launch list of the same command + full captures of the five hot kernels (one ncu call group)
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
B="python bench.py --steps 3 --warmup 3 --eager --no-cpu-baseline --no-ref-gpu --no-e2e"
$B > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 90 -c 30 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launches.log 2>&1
for spec in "warp_bwd:warp_bwd" "warp_fwd:warp_fwd" "corr_fwd:corr_tc_fwd" "corr_bwd:corr_tc_bwd" "ds6x8:downsample"; do
  op=${spec%%:*}; rx=${spec##*:}
  P="python tools/prof_op.py $op 32 tf32"
  $P > gpurun_out/plain_$op.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:$rx -s 2 -c 1 -f -o gpurun_out/prof_$op $P > gpurun_out/ncu_$op.log 2>&1
  tail -2 gpurun_out/ncu_$op.log
done
cat gpurun_out/bench_default.json; ls -la gpurun_out
