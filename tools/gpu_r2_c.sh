#!/bin/bash
# round 2, visit C: tests after the mapping change + attribution of the step time
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_fullsize.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
(timeout 300 python tools/time_step.py 1048576:3 65536:40 4096:100
NB_LIB_PATH=build/libnb_nonear.so timeout 300 python tools/time_step.py 1048576:3 65536:40
timeout 200 python tools/time_step.py 1048576:2 65536:40 2
NB_FORCE_R=2 timeout 200 python tools/time_step.py 1048576:2
) > gpurun_out/time_step.txt 2>&1
cat gpurun_out/time_step.txt
