#!/bin/bash
# round 2, trip a: split-tf32 correlation -- parity + per-op times
mkdir -p gpurun_out
for t in test_gpu_tc test_reference_pin test_gpu_fused test_gpu_graphs test_gpu_downsample_multi; do
  timeout 900 python -m pytest tests/$t.py -q -m gpu > gpurun_out/r2a_$t.log 2>&1; echo "exit $?" >> gpurun_out/r2a_$t.log
  tail -4 gpurun_out/r2a_$t.log
done
timeout 300 python tools/time_ops.py 32 tf32x3 corr > gpurun_out/r2a_time_x3.jsonl 2>&1; cat gpurun_out/r2a_time_x3.jsonl
timeout 300 python tools/time_ops.py 32 tf32 corr > gpurun_out/r2a_time_tf32.jsonl 2>&1; cat gpurun_out/r2a_time_tf32.jsonl
timeout 300 python tools/time_ops.py 8 tf32x3 corr > gpurun_out/r2a_time_x3_n8.jsonl 2>&1; cat gpurun_out/r2a_time_x3_n8.jsonl
timeout 300 python tools/time_ops.py 32 tf32 down > gpurun_out/r2a_time_ds.jsonl 2>&1; cat gpurun_out/r2a_time_ds.jsonl
timeout 300 python tools/time_ops.py 8 tf32 down > gpurun_out/r2a_time_ds_n8.jsonl 2>&1; cat gpurun_out/r2a_time_ds_n8.jsonl
