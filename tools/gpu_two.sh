#!/bin/bash
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_multi.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/pytest_two.log 2>&1; tail -3 gpurun_out/pytest_two.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29702 bench.py --gpus 2 --steps 5 --warmup 3 --cpu-sample 0 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench2 exit $?"
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench1 exit $?"
python - <<'PY'
import json
for f in ("bench_n2","bench"):
    d=json.load(open(f"gpurun_out/{f}.json")); print(f, d["value"], d["ms_per_step"], d["e2e"], d.get("general_mass"), d.get("planar_2d"))
PY
