#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_tc.py tests/test_gpu_graphs.py tests/test_gpu_parity.py tests/test_reference_pin.py -q -m gpu -x > gpurun_out/r2k.log 2>&1; echo "exit $?" >> gpurun_out/r2k.log
tail -6 gpurun_out/r2k.log
timeout 300 python tools/time_ops.py 32 tf32x3 corr_bwd
timeout 300 python tools/time_ops.py 32 tf32 corr_bwd
