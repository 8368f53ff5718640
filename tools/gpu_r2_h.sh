#!/bin/bash
# round 2, visit H: team-refactored host code on one GPU + kernel variants (H trimmed prologue, I +3 CTAs/SM, J +flush per group, K flush per group)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_fullsize.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|exit|Error" gpurun_out/pytest_gpu.log | head -40
(for rep in 1 2; do
NB_LIB_PATH=build/libnb_r1.so timeout 300 python tools/time_step.py 1048576:6 65536:60
timeout 300 python tools/time_step.py 1048576:6 65536:60
for v in I J K; do
NB_LIB_PATH=build/libnb_$v.so timeout 300 python tools/time_step.py 1048576:6 65536:60
done; done
) > gpurun_out/time_variants.txt 2>&1
cat gpurun_out/time_variants.txt
NB_LIB_PATH=build/libnb_J.so timeout 600 python tests/tools/evolved_accuracy.py disk4096 disk3d4096 plummer16384 plummer65536 > gpurun_out/evolved_accuracy_J.txt 2>&1
cat gpurun_out/evolved_accuracy_J.txt
