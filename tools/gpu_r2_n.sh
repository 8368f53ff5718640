#!/bin/bash
# round 2, visit N (2 GPUs): multi-GPU tests again after the tolerance / scenario fixes; config 5 at 2 and 1 GPUs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_dropin.py -m gpu -q -rs > gpurun_out/pytest_multi_n2.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi_n2.log
tail -n 12 gpurun_out/pytest_multi_n2.log
NB_DROPIN_DEVICES=0,1 NB_DROPIN_ONLY=F timeout 300 oracle/_ref/dropin_check > gpurun_out/dropin_multi.txt 2>&1; cat gpurun_out/dropin_multi.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29602 bench.py --config 5 --gpus 2 --steps 3 --warmup 3 --cpu-sample 0 > gpurun_out/config5_n2.json 2> gpurun_out/config5_n2.err; echo "config5 n=2 exit $?"; cat gpurun_out/config5_n2.json; tail -n 2 gpurun_out/config5_n2.err
