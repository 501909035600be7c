#!/bin/bash
# first GPU call of the session: tests, bench (fp32 + tf32), launch list under ncu
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
python tools/check_tc.py tf32 > gpurun_out/check_tc.log 2>&1
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_fp32.json 2> gpurun_out/bench_fp32.err
python bench.py --steps 10 --warmup 3 --precision tf32 --no-cpu-baseline --no-ref-gpu > gpurun_out/bench_tf32.json 2> gpurun_out/bench_tf32.err
python bench.py --steps 2 --warmup 3 --precision tf32 --no-cpu-baseline --no-ref-gpu --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_tf32.csv \
  python bench.py --steps 2 --warmup 3 --precision tf32 --no-cpu-baseline --no-ref-gpu --no-e2e > gpurun_out/ncu_launches.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/check_tc.log | tail -20; cat gpurun_out/bench_fp32.json gpurun_out/bench_tf32.json
