#!/bin/bash
# usage: gpu_scale.sh N  -- both bench arms at N GPUs, launched the way the driver does
N=$1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/scale_ref_$N.json 2> gpurun_out/scale_ref_$N.err
tail -c 300 gpurun_out/scale_ref_$N.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err
tail -c 1500 gpurun_out/scale_$N.json
