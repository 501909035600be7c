#!/bin/bash
# session check: gpu tests, per-op device times (CUDA graphs), one bench line
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.log 2>&1
nproc >> gpurun_out/smi.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python tools/time_ops.py 32 tf32 > gpurun_out/time_ops_tf32.log 2>&1
timeout 300 python tools/time_ops.py 8 tf32 warp,down > gpurun_out/time_ops_n8.log 2>&1
timeout 600 python bench.py --precision tf32 > gpurun_out/bench_tf32.json 2> gpurun_out/bench_tf32.err
tail -5 gpurun_out/pytest_gpu.log; cat gpurun_out/time_ops_tf32.log gpurun_out/time_ops_n8.log; cat gpurun_out/bench_tf32.json; tail -3 gpurun_out/bench_tf32.err
