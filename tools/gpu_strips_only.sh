#!/bin/bash
N=$1
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --config strips4k --steps 50 --warmup 5 > gpurun_out/strips4k_$N.json 2> gpurun_out/strips4k_$N.err
tail -c 1200 gpurun_out/strips4k_$N.json; echo; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/strips4k_$N.err | tail -5
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus $N --config strips4k --precision tf32 --steps 50 --warmup 5 > gpurun_out/strips4k_tf32_$N.json 2>> gpurun_out/strips4k_$N.err
tail -c 700 gpurun_out/strips4k_tf32_$N.json; echo
