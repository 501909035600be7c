#!/bin/bash
# last check of the session: all gpu tests, smoke, bench (both arms), per-op times
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/s24y_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/s24y_pytest.log
tail -3 gpurun_out/s24y_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/s24y_bench.json 2> gpurun_out/s24y_bench.err; tail -c 300 gpurun_out/s24y_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/s24y_bench_ref.json 2> gpurun_out/s24y_bench_ref.err
python tools/time_ops.py 32 tf32 > gpurun_out/s24y_time_ops.jsonl 2>&1
python tools/time_ops.py 8 tf32 > gpurun_out/s24y_time_ops_n8.jsonl 2>&1
python tools/timeline_ds.py 32 > gpurun_out/s24y_timeline_ds.txt 2>&1
tail -4 gpurun_out/s24y_time_ops_n8.jsonl
