#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python tools/tune.py 1048576 2 > gpurun_out/tune20.txt 2>&1
tail -4 gpurun_out/pytest_gpu.log; cat gpurun_out/tune20.txt
