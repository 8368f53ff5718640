#!/bin/bash
# round 2, visit P (8 GPUs): multi-GPU tests at 8 and 3 ranks, bench at 8 and 4, config 5 at 8 and 4, binding on 8 GPUs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_dropin.py -m gpu -q -rs > gpurun_out/pytest_multi_n8.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi_n8.log
tail -n 8 gpurun_out/pytest_multi_n8.log
NB_DROPIN_DEVICES=0,1,2,3,4,5,6,7 NB_DROPIN_ONLY=F timeout 300 oracle/_ref/dropin_check > gpurun_out/dropin_multi_n8.txt 2>&1; cat gpurun_out/dropin_multi_n8.txt
for n in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 5 --warmup 3 --cpu-sample 0 --no-tree > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  echo "bench n=$n exit $?"; cut -c1-250 gpurun_out/bench_n$n.json; tail -n 2 gpurun_out/bench_n$n.err
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --config 5 --gpus $n --steps 5 --warmup 3 --cpu-sample 0 > gpurun_out/config5_n$n.json 2> gpurun_out/config5_n$n.err
  echo "config5 n=$n exit $?"; cut -c1-250 gpurun_out/config5_n$n.json; tail -n 2 gpurun_out/config5_n$n.err
done
