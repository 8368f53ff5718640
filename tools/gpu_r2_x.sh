#!/bin/bash
# round 2, visit X (2 GPUs): ownership along one curve: multi-GPU tests, the binding on 2 GPUs, bench at 2 GPUs with and without it
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_dropin.py -m gpu -q -rs -x > gpurun_out/pytest_multi_x.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi_x.log
tail -n 30 gpurun_out/pytest_multi_x.log
for cp in 1 0; do
  NB_CURVE_PARTITION=$cp timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29510+cp)) bench.py --gpus 2 --steps 5 --warmup 3 --cpu-sample 0 --no-tree > gpurun_out/bench_n2_cp$cp.json 2> gpurun_out/bench_n2_cp$cp.err
  echo "bench cp=$cp exit $?"; python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_n2_cp$cp.json"))
    print({k: d.get(k) for k in ("ms_per_step", "value", "e2e", "near_path", "parity", "exchange_overlap")})
except Exception as e:
    print("no json", e)
PY
  tail -n 3 gpurun_out/bench_n2_cp$cp.err
done
