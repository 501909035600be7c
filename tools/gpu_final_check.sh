#!/bin/bash
# the driver's round-end sequence on one GPU: all gpu tests, smoke(), both bench arms
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/final_pytest.log; tail -3 gpurun_out/final_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/final_bench_ref.json 2>gpurun_out/final_bench_ref.err; tail -c 200 gpurun_out/final_bench_ref.json; echo
python bench.py --steps 20 --warmup 5 > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; python - <<'PY'
import json
l = json.loads([x for x in open("gpurun_out/final_bench.json") if x.startswith("{")][-1])
print({k: l[k] for k in ("value", "ms_per_step", "dtype", "gpu_launches")}, l["roofline"]["kernel"], l["roofline"]["frac"], l["roofline"]["traffic"], "e2e", round(l["e2e"]["value"]))
PY
