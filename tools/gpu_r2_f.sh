#!/bin/bash
# round 2, visit F: tests + diag debug + variants after restoring uniform-register addressing in the hot loop
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_fullsize.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|exit|Error" gpurun_out/pytest_gpu.log | head -40
timeout 300 python tools/debug_diag.py > gpurun_out/debug_diag.txt 2>&1; cat gpurun_out/debug_diag.txt
(for rep in 1 2; do
NB_LIB_PATH=build/libnb_r1.so timeout 300 python tools/time_step.py 1048576:6 65536:60
timeout 300 python tools/time_step.py 1048576:6 65536:60
for v in C D G; do
NB_LIB_PATH=build/libnb_$v.so timeout 300 python tools/time_step.py 1048576:6 65536:60
done; done
timeout 300 python tools/time_step.py 262144:20 16384:100 4096:100
timeout 200 python tools/time_step.py 1048576:4 65536:60 2
) > gpurun_out/time_variants.txt 2>&1
cat gpurun_out/time_variants.txt
