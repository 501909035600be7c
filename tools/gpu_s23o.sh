#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tc.py tests/test_reference_pin.py tests/test_gpu_strips.py tests/test_gpu_fused.py -x -q > gpurun_out/s23o_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/s23o_pytest.log
tail -15 gpurun_out/s23o_pytest.log
timeout 300 python tools/time_ops.py 32 tf32 corr_fwd 2>&1 | tail -1
FN2_PAIR_V2=1 timeout 300 python tools/time_ops.py 32 tf32 corr_fwd 2>&1 | tail -1
for n in 8 64; do timeout 300 python tools/time_ops.py $n tf32 corr_fwd 2>&1 | tail -1; done
