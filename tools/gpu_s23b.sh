#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_downsample_multi.py -x -q > gpurun_out/s23b_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/s23b_pytest.log
tail -30 gpurun_out/s23b_pytest.log
timeout 300 python tools/time_ops.py 32 tf32 down > gpurun_out/s23b_time_ds.jsonl 2>&1; cat gpurun_out/s23b_time_ds.jsonl
timeout 300 python tools/time_ops.py 8 tf32 down > gpurun_out/s23b_time_ds8.jsonl 2>&1; tail -1 gpurun_out/s23b_time_ds8.jsonl
bash tools/gpu_prof.sh ds_multi downsample_multi ds_multi 32
