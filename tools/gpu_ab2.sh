#!/bin/bash
echo "== default"; timeout 300 python tools/time_warp.py 32
for l in "$@"; do echo "== $l"; FN2_OPS_LIBRARY=$PWD/$l timeout 300 python tools/time_warp.py 32; done
