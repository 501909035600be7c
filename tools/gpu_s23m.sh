#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_reference_pin.py tests/test_golden.py -x -q -k "warp or golden or pin" 2>&1 | tail -4
timeout 300 python tools/time_ops.py 32 tf32 warp_bwd 2>&1 | tail -1
FN2_WARP_BWD_SPLIT=1 timeout 300 python tools/time_ops.py 32 tf32 warp_bwd 2>&1 | tail -1
