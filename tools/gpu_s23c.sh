#!/bin/bash
mkdir -p gpurun_out
for v in 0 1 2 3 4; do echo "variant $v"; FN2_DS_VARIANT=$v timeout 300 python tools/time_ops.py 32 tf32 down 2>&1 | tail -1; done
