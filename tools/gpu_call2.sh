#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
for b in fp64mix pipes mix inner; do timeout 120 tools/ubench/$b > gpurun_out/ubench_$b.txt 2>&1; done
tail -5 gpurun_out/pytest_gpu.log; cat gpurun_out/ubench_fp64mix.txt
