#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tc.py -q -m gpu -x -k "2_to_the_31" 2>&1 | tail -8
