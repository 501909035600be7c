#!/bin/bash
# round 2 evidence: all gpu tests, bench (both arms), per-op times, launch list, full ncu captures of every hot kernel
# usage: tools/gpu_r2_evidence.sh <tag>     -> gpurun_out/<tag>_*; then python tools/make_traffic.py <tag>
T=${1:-r02}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest_gpu.log
tail -3 gpurun_out/${T}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_ref.err
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; tail -c 300 gpurun_out/${T}_bench.json; tail -3 gpurun_out/${T}_bench.err
python bench.py --precision tf32 --no-extras --no-cpu-baseline --no-ref-gpu > gpurun_out/${T}_bench_tf32.json 2>> gpurun_out/${T}_bench.err
python tools/time_ops.py 32 tf32x3 > gpurun_out/${T}_time_ops_x3.jsonl 2>&1
python tools/time_ops.py 32 tf32 corr > gpurun_out/${T}_time_ops_tf32.jsonl 2>&1
python tools/time_ops.py 8 tf32x3 > gpurun_out/${T}_time_ops_x3_n8.jsonl 2>&1
B="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --no-ref-gpu --no-e2e"
$B > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 60 --csv --log-file gpurun_out/${T}_launches.csv $B > gpurun_out/ncu_launches.log 2>&1
bash tools/gpu_prof.sh x3_fwd corr_tc_fwd_x3_pair corr_fwd 32 tf32x3
bash tools/gpu_prof.sh x3_bwd corr_tc_bwd_x3_pair corr_bwd 32 tf32x3
bash tools/gpu_prof.sh corr_fwd corr_tc_fwd_pair corr_fwd 32 tf32
bash tools/gpu_prof.sh corr_bwd corr_tc_bwd corr_bwd 32 tf32
bash tools/gpu_prof.sh warp_fwd warp_tile_gather warp_fwd 32
bash tools/gpu_prof.sh warp_bwd warp_tile_bwd warp_bwd 32
bash tools/gpu_prof.sh ds_band ms_band ds_multi 32
bash tools/gpu_prof.sh ds_combine ms_combine ds_multi 32
bash tools/gpu_prof.sh ds_tables ms_tables ds_multi 32
ls gpurun_out | wc -l
