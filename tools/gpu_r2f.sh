#!/bin/bash
mkdir -p gpurun_out
for t in test_gpu_parity test_reference_pin test_gpu_fixtures test_gpu_fused; do
  timeout 900 python -m pytest tests/$t.py -q -m gpu -x -k "warp or fixtures or sample" > gpurun_out/r2f_$t.log 2>&1; echo "exit $?" >> gpurun_out/r2f_$t.log
  tail -6 gpurun_out/r2f_$t.log
done
timeout 300 python tools/time_ops.py 32 tf32 warp > gpurun_out/r2f_time_warp.jsonl 2>&1; cat gpurun_out/r2f_time_warp.jsonl
timeout 300 python tools/time_ops.py 8 tf32 warp > gpurun_out/r2f_time_warp_n8.jsonl 2>&1; cat gpurun_out/r2f_time_warp_n8.jsonl
