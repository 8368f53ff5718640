#!/bin/bash
# round 2, visit Q (2 GPUs): final library — full single-GPU suite, small-world frame times, bench at 2 GPUs (re-upload path under NCCL)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|exit" gpurun_out/pytest_gpu.log | head -20
timeout 300 python tools/time_configs.py small > gpurun_out/time_small.txt 2>&1; cat gpurun_out/time_small.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 bench.py --gpus 2 --steps 5 --warmup 3 --cpu-sample 0 --no-tree > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n=2 exit $?"; cut -c1-200 gpurun_out/bench_n2.json; tail -n 2 gpurun_out/bench_n2.err
