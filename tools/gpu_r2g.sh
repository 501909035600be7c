#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_fused.py tests/test_reference_pin.py -q -m gpu -x > gpurun_out/r2g_tc.log 2>&1; echo "exit $?" >> gpurun_out/r2g_tc.log
tail -15 gpurun_out/r2g_tc.log
timeout 300 python tools/time_ops.py 32 tf32x3 corr > gpurun_out/r2g_time_x3.jsonl 2>&1; cat gpurun_out/r2g_time_x3.jsonl
timeout 300 python tools/time_ops.py 8 tf32x3 corr_fwd > gpurun_out/r2g_time_x3_n8.jsonl 2>&1; cat gpurun_out/r2g_time_x3_n8.jsonl
FN2_NO_PAIR=1 timeout 300 python tools/time_ops.py 32 tf32x3 corr_fwd
python tools/dbg_smoke.py 2>&1 | tail -12
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
