#!/bin/bash
# tcgen05 backward bring-up: error statistics for both descriptor variants, then timing
mkdir -p gpurun_out
for v in 0 1; do
  echo "=== FN2_TC_BWD_VARIANT=$v" >> gpurun_out/check_tc_bwd.log
  FN2_TC_BWD_VARIANT=$v timeout 300 python tools/check_tc.py tf32 >> gpurun_out/check_tc_bwd.log 2>&1
  echo "rc=$?" >> gpurun_out/check_tc_bwd.log
done
cat gpurun_out/check_tc_bwd.log
