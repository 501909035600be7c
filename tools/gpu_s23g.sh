#!/bin/bash
for d in 0 1 2; do echo "dbg $d"; FN2_DBG=$d timeout 300 python tools/time_ops.py 32 tf32 corr_fwd 2>&1 | tail -1; done
