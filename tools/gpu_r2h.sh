#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python -m pytest tests/test_gpu_tc.py -q -m gpu -x -k "fresh_thread" 2>&1 | tail -3
bash tools/gpu_prof.sh x3_pair corr_tc_fwd_x3_pair corr_fwd 32 tf32x3
