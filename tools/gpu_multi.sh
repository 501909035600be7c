#!/bin/bash
# usage: gpu_multi.sh N  -- multi-GPU evidence at N GPUs: NCCL strip parity + timing, pytest multi-GPU tests, strips4k
# bench (strong scaling, NCCL send/recv inside the timed region), both default bench arms as the driver launches them
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for p in tf32 tf32x3; do
  timeout 300 $TR --master-port 29531 tools/strips_nccl.py $p > gpurun_out/multi_strips_${p}_$N.json 2> gpurun_out/multi_strips_${p}_$N.err
  tail -c 700 gpurun_out/multi_strips_${p}_$N.json; echo
done
if [ "$N" = "2" ]; then timeout 600 python -m pytest tests/test_gpu_strips.py tests/test_gpu_tc.py -q -m gpu -k "nccl or two_devices" 2>&1 | tail -3; fi
timeout 300 $TR --master-port 29532 bench.py --gpus $N --config strips4k --steps 50 --warmup 5 > gpurun_out/multi_strips4k_$N.json 2> gpurun_out/multi_strips4k_$N.err
tail -c 900 gpurun_out/multi_strips4k_$N.json; echo
timeout 300 $TR --master-port 29533 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/multi_ref_$N.json 2> gpurun_out/multi_ref_$N.err
timeout 600 $TR --master-port 29534 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/multi_bench_$N.json 2> gpurun_out/multi_bench_$N.err
python - <<PY
import json
try:
    l = json.loads([x for x in open("gpurun_out/multi_bench_$N.json").read().splitlines() if x.startswith("{")][-1])
    print({k: l[k] for k in ("value", "ms_per_step", "n_gpus")}, "e2e", l["e2e"]["value"], "ceiling", l["e2e_copy_ceiling"]["value"])
    print("strips4k", json.dumps(l.get("strips4k"))[:700])
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/multi_bench_$N.err").read()[-1500:])
PY
