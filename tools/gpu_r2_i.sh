#!/bin/bash
# round 2, visit I: device merge replay + near-path variants
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_fullsize.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|exit|Error" gpurun_out/pytest_gpu.log | head -40
timeout 900 python tests/tools/time_merge.py big > gpurun_out/time_merge.txt 2>&1; cat gpurun_out/time_merge.txt
(for rep in 1 2; do
timeout 300 python tools/time_step.py 1048576:6 65536:60 16384:100
for v in L M N; do
NB_LIB_PATH=build/libnb_$v.so timeout 300 python tools/time_step.py 1048576:6 65536:60 16384:100
done; done
) > gpurun_out/time_variants.txt 2>&1
cat gpurun_out/time_variants.txt
