#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_fused.py tests/test_gpu_graphs.py -q -m gpu -x > gpurun_out/r2j_tc.log 2>&1; echo "exit $?" >> gpurun_out/r2j_tc.log
tail -4 gpurun_out/r2j_tc.log
timeout 300 python tools/time_ops.py 32 tf32x3 corr
timeout 300 python tools/time_ops.py 8 tf32x3 corr
