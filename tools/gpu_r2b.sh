#!/bin/bash
# round 2, trip b: fixes after trip a (x3 accumulators, 2-row x3 forward, downsample tables/NaN rule), strips, bench
mkdir -p gpurun_out
for t in test_gpu_tc test_gpu_downsample_multi test_gpu_strips test_reference_pin test_gpu_graphs test_gpu_fused; do
  timeout 900 python -m pytest tests/$t.py -q -m gpu > gpurun_out/r2b_$t.log 2>&1; echo "exit $?" >> gpurun_out/r2b_$t.log
  tail -4 gpurun_out/r2b_$t.log
done
timeout 300 python tools/time_ops.py 32 tf32x3 corr > gpurun_out/r2b_time_x3.jsonl 2>&1; cat gpurun_out/r2b_time_x3.jsonl
timeout 300 python tools/time_ops.py 32 tf32 downsample_multi,multiscale > gpurun_out/r2b_time_ds.jsonl 2>&1; cat gpurun_out/r2b_time_ds.jsonl
timeout 300 python tools/time_ops.py 8 tf32 downsample_multi,multiscale > gpurun_out/r2b_time_ds_n8.jsonl 2>&1; cat gpurun_out/r2b_time_ds_n8.jsonl
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; tail -c 3000 gpurun_out/r2b_bench.json; tail -5 gpurun_out/r2b_bench.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2b_bench_ref.json 2> gpurun_out/r2b_bench_ref.err; tail -c 600 gpurun_out/r2b_bench_ref.json
python tools/prof_multi.py 32 > gpurun_out/plain_multi.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2b_launches.csv python tools/prof_multi.py 32 > gpurun_out/ncu_multi.log 2>&1
grep -v "^==" gpurun_out/r2b_launches.csv | awk -F'","' '{print $5, $NF}' | tail -40
