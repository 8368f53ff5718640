#!/bin/bash
# round 2: config 5 (2^22-body two-galaxy collision) with ownership along the curve; usage: bash tools/gpu_r2_c5.sh <n_gpus>
N=${1:-8}
mkdir -p gpurun_out
timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --config 5 --gpus $N --steps 5 --warmup 3 --cpu-sample 0 > gpurun_out/config5_n$N.json 2> gpurun_out/config5_n$N.err
echo "config5 n=$N exit $?"; cat gpurun_out/config5_n$N.json; tail -n 2 gpurun_out/config5_n$N.err
