#!/bin/bash
mkdir -p gpurun_out
for s in 8 2 0.5; do timeout 60 tools/mb/mb_scatter $s; done > gpurun_out/mb_scatter.log 2>&1; cat gpurun_out/mb_scatter.log
