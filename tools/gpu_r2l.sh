#!/bin/bash
mkdir -p gpurun_out
bash tools/gpu_prof.sh x3_bwd_pair corr_tc_bwd_x3_pair corr_bwd 32 tf32x3
