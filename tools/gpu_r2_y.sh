#!/bin/bash
# round 2, visit Y (8 GPUs): ownership along one curve at 8 GPUs: the 8- and 3-GPU test cases, binding on 8 GPUs, bench at 8 and 4
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_dropin.py -m gpu -q -rs -k "8-9000 or 3-5000-1 or 3-1500 or 3-1700" > gpurun_out/pytest_multi_y.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi_y.log
tail -n 12 gpurun_out/pytest_multi_y.log
NB_DROPIN_DEVICES=0,1,2,3,4,5,6,7 NB_DROPIN_ONLY=F timeout 200 oracle/_ref/dropin_check > gpurun_out/dropin_multi_n8.txt 2>&1; tail -n 4 gpurun_out/dropin_multi_n8.txt
for n in 8 4; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520+n)) bench.py --gpus $n --steps 5 --warmup 3 --cpu-sample 0 --no-tree > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  echo "bench n=$n exit $?"; python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_n$n.json"))
    print({k: d.get(k) for k in ("ms_per_step", "value", "e2e", "near_path", "exchange_overlap")}); print(d["parity"]["ok"], d["parity"]["max_rel_err"], d["state_digest"])
except Exception as e:
    print("no json", e)
PY
  tail -n 2 gpurun_out/bench_n$n.err
done
