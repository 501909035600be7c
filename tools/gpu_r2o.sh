#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_reference_pin.py tests/test_gpu_fixtures.py tests/test_gpu_fused.py tests/test_golden.py -q -m gpu -x -k "warp or fixtures or sample or golden" 2>&1 | tail -4
python tools/time_warp.py 32; python tools/time_warp.py 8
