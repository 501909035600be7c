#!/bin/bash
# evidence of the session: all gpu tests, bench (both arms), launch list, per-op times, full ncu captures
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/s24z_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/s24z_pytest.log
tail -3 gpurun_out/s24z_pytest.log
python bench.py > gpurun_out/s24z_bench.json 2> gpurun_out/s24z_bench.err; tail -c 400 gpurun_out/s24z_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/s24z_bench_ref.json 2> gpurun_out/s24z_bench_ref.err
python tools/time_ops.py 32 tf32 > gpurun_out/s24z_time_ops.jsonl 2>&1
B="python bench.py --steps 2 --warmup 3 --precision tf32 --no-cpu-baseline --no-ref-gpu --no-e2e"
$B > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 80 -c 40 --csv --log-file gpurun_out/s24z_launches_tf32.csv $B > gpurun_out/ncu_launches.log 2>&1
bash tools/gpu_prof.sh corr_fwd corr_tc_fwd_pair corr_fwd 32
bash tools/gpu_prof.sh corr_bwd corr_tc_bwd corr_bwd 32
bash tools/gpu_prof.sh warp_fwd warp_tile_gather warp_fwd 32
bash tools/gpu_prof.sh warp_bwd warp_tile_bwd warp_bwd 32
bash tools/gpu_prof.sh ds_multi downsample_multi ds_multi 32
ls gpurun_out | wc -l
