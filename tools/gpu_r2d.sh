#!/bin/bash
# round 2, trip d: downsample (trimmed runs, combine table), scatter microbenchmark
mkdir -p gpurun_out
for t in test_gpu_downsample_multi test_gpu_fixtures; do
  timeout 900 python -m pytest tests/$t.py -q -m gpu -x > gpurun_out/r2d_$t.log 2>&1; echo "exit $?" >> gpurun_out/r2d_$t.log
  tail -15 gpurun_out/r2d_$t.log
done
timeout 300 python tools/time_ops.py 32 tf32 downsample_multi,multiscale > gpurun_out/r2d_time_ds.jsonl 2>&1; cat gpurun_out/r2d_time_ds.jsonl
timeout 300 python tools/time_ops.py 8 tf32 downsample_multi,multiscale > gpurun_out/r2d_time_ds_n8.jsonl 2>&1; cat gpurun_out/r2d_time_ds_n8.jsonl
python tools/prof_multi.py 32 ds_multi epe > gpurun_out/plain_multi.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2d_launches.csv python tools/prof_multi.py 32 ds_multi epe > gpurun_out/ncu_multi.log 2>&1
grep "fn2::" gpurun_out/r2d_launches.csv | awk -F'","' '{print $5, $NF}' | tail -6
timeout 120 tools/mb/mb_scatter 8 > gpurun_out/r2d_mb_scatter.log 2>&1; timeout 120 tools/mb/mb_scatter 2 >> gpurun_out/r2d_mb_scatter.log 2>&1; cat gpurun_out/r2d_mb_scatter.log
