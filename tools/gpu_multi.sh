#!/bin/bash
# One visit to an 8-GPU box: the driver's scaling bench at N = 8, 4, 2, the multi-rank parity tests, config 5.
mkdir -p gpurun_out
nvidia-smi -L | wc -l > gpurun_out/ngpu.txt
for n in 8 4 2; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 5 --warmup 3 --cpu-sample 0 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  echo "bench n=$n exit $?"
done
timeout 500 python -m pytest tests/test_gpu_multi.py -m gpu -q > gpurun_out/pytest_multi.log 2>&1; tail -3 gpurun_out/pytest_multi.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29608 tools/config5.py 4194304 10 > gpurun_out/config5_n8.jsonl 2> gpurun_out/config5_n8.err
cat gpurun_out/config5_n8.jsonl
python - <<'PY'
import json
for n in (8, 4, 2):
    try:
        d = json.load(open(f"gpurun_out/bench_n{n}.json")); print(n, d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["h2d_bytes_per_step"], d["clocks"])
    except Exception as e: print(n, "failed", e)
PY
