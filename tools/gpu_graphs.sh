#!/bin/bash
N=${1:-1}
mkdir -p gpurun_out
if [ "$N" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"; fi
timeout 600 $L bench.py --gpus $N --config flownetc_train --steps 10 --warmup 3 > gpurun_out/graphs_train_$N.json 2> gpurun_out/graphs_train_$N.err; grep "^{" gpurun_out/graphs_train_$N.json | tail -1 | cut -c1-400; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/graphs_train_$N.err | tail -4
timeout 600 $L bench.py --gpus $N --config flownet2_infer --steps 10 --warmup 3 > gpurun_out/graphs_infer_$N.json 2> gpurun_out/graphs_infer_$N.err; grep "^{" gpurun_out/graphs_infer_$N.json | tail -1 | cut -c1-400; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/graphs_infer_$N.err | tail -4
timeout 600 $L bench.py --gpus $N --config flownetc_train --precision tf32 --steps 10 --warmup 3 2>/dev/null | grep "^{" | tail -1 | cut -c1-200
