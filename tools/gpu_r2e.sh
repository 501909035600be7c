#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fixtures.py -q -m gpu > gpurun_out/r2e_fixtures.log 2>&1; tail -5 gpurun_out/r2e_fixtures.log
timeout 60 tools/mb/mb_scatter 8 > gpurun_out/r2e_mb_scatter.log 2>&1; timeout 60 tools/mb/mb_scatter 2 >> gpurun_out/r2e_mb_scatter.log 2>&1; cat gpurun_out/r2e_mb_scatter.log
