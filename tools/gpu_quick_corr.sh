#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_strips.py tests/test_gpu_tc.py -x -q > gpurun_out/pytest_corr.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_corr.log
timeout 300 python tools/time_ops.py 32 tf32 corr > gpurun_out/time_corr.log 2>&1
timeout 300 python tools/time_4k.py tf32 > gpurun_out/time_4k.log 2>&1
tail -4 gpurun_out/pytest_corr.log; cat gpurun_out/time_corr.log; head -2 gpurun_out/time_4k.log
