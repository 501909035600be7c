#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_fused.py -q -m gpu -x > gpurun_out/r2i_tc.log 2>&1; echo "exit $?" >> gpurun_out/r2i_tc.log
tail -4 gpurun_out/r2i_tc.log
timeout 300 python tools/time_ops.py 32 tf32x3 corr_fwd
timeout 300 python tools/time_ops.py 8 tf32x3 corr_fwd
bash tools/gpu_prof.sh x3_pair corr_tc_fwd_x3_pair corr_fwd 32 tf32x3
