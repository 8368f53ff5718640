#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/config5_n8.jsonl
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29608 tools/config5.py 4194304 10 >> gpurun_out/config5_n8.jsonl 2> gpurun_out/config5_n8.err
echo "c5 n=8 exit $?"
timeout 500 python -m pytest tests/test_gpu_multi.py -m gpu -q > gpurun_out/pytest_multi.log 2>&1; tail -3 gpurun_out/pytest_multi.log
cat gpurun_out/config5_n8.jsonl
