#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_reference_pin.py tests/test_golden.py -x -q -k "warp or golden" > gpurun_out/pytest_warp.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_warp.log
timeout 300 python tools/time_ops.py 32 tf32 warp > gpurun_out/time_warp.log 2>&1
tail -6 gpurun_out/pytest_warp.log; cat gpurun_out/time_warp.log
