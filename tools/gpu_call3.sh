#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 --precision tf32 --no-ref-gpu > gpurun_out/bench_tf32.json 2> gpurun_out/bench_tf32.err
tail -5 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_tf32.json; tail -3 gpurun_out/bench_tf32.err
