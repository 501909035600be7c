#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -k "downsample or warp" > gpurun_out/pytest_ds.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_ds.log
tail -15 gpurun_out/pytest_ds.log
python tools/time_ops.py 32 tf32 warp,down > gpurun_out/time_ds.log 2>&1; cat gpurun_out/time_ds.log
