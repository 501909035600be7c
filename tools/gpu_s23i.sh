#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_graphs.py -x -q -s > gpurun_out/s23i_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/s23i_pytest.log
tail -30 gpurun_out/s23i_pytest.log
