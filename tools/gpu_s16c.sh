#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
timeout 600 python bench.py --eager --steps 200 --no-e2e --no-cpu-baseline --no-ref-gpu > gpurun_out/bench_eager.json 2> gpurun_out/bench_eager.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
cat gpurun_out/bench_default.json; tail -5 gpurun_out/bench_default.err; cat gpurun_out/bench_eager.json gpurun_out/bench_ref.json
