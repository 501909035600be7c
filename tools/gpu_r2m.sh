#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_strips.py tests/test_gpu_fused.py -q -m gpu -x > gpurun_out/r2m.log 2>&1; echo "exit $?" >> gpurun_out/r2m.log
tail -6 gpurun_out/r2m.log
python tools/time_4k.py tf32x3 2>&1 | grep -E "1of1|1of4"
timeout 300 python tools/time_ops.py 32 tf32x3 corr_fwd
