#!/bin/bash
# round 2, visit M (2 GPUs): both multi-GPU transports against the single-GPU run, the flecs binding on 2 GPUs, bench at N=2
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_collision_merge.py -m gpu -q > gpurun_out/pytest_merge.log 2>&1; tail -n 2 gpurun_out/pytest_merge.log
timeout 100 python tests/tools/time_merge.py > gpurun_out/time_merge.txt 2>&1; cat gpurun_out/time_merge.txt
bash tools/gpu_r2_multi.sh 2
