#!/bin/bash
# round 2, visit D: A/B of the step time against the round-1 kernels in one visit; Hilbert order, near factor 0.25
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_fullsize.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|exit|AssertionError" gpurun_out/pytest_gpu.log | head -40
(for rep in 1 2; do
NB_LIB_PATH=build/libnb_r1.so timeout 300 python tools/time_step.py 1048576:8 65536:60
timeout 300 python tools/time_step.py 1048576:8 65536:60
NB_LIB_PATH=build/libnb_once.so timeout 300 python tools/time_step.py 1048576:8 65536:60
NB_LIB_PATH=build/libnb_nonear.so timeout 300 python tools/time_step.py 1048576:8 65536:60
done
timeout 300 python tools/time_step.py 262144:20 16384:100 4096:100
timeout 200 python tools/time_step.py 1048576:4 65536:60 2
) > gpurun_out/time_step.txt 2>&1
cat gpurun_out/time_step.txt
