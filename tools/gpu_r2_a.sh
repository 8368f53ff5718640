#!/bin/bash
# round 2, visit A: baseline state of the tree + evolved-state accuracy of the FP32 path
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python tests/tools/evolved_accuracy.py disk4096 disk3d4096 plummer16384 plummer65536 > gpurun_out/evolved_accuracy.txt 2>&1
cat gpurun_out/evolved_accuracy.txt
timeout 300 python bench.py --steps 3 --warmup 3 --no-tree --cpu-sample 0 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json
