#!/bin/bash
N=$1
mkdir -p gpurun_out
for mr in 0 100000; do for p in tf32x3 tf32; do
FN2_STRIPS_OVERLAP_MIN_ROWS=$mr timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --config strips4k --precision $p --steps 50 --warmup 5 2>/dev/null | grep "^{" | tail -1 | python -c "
import sys, json
l = json.loads(sys.stdin.read()); s = l['strips4k']
print('min_rows', $mr, s['precision'], 'gpus', s['n_gpus'], 'graph ms', s['ms_per_step'], 'eager ms', s['ms_per_step_eager'], 'exchange', s['collective']['exchange_alone_ms'])"
done; done
