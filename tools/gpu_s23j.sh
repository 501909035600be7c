#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_graphs.py tests/test_gpu_tc.py tests/test_gpu_parity.py -x -q > gpurun_out/s23j_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/s23j_pytest.log
tail -25 gpurun_out/s23j_pytest.log
