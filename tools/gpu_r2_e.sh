#!/bin/bash
# round 2, visit E: tests + kernel variant A/B (A baseline, B int flush, C 3 CTAs/SM, D 3 stages, E B+C, F unroll 16)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_fullsize.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|exit|Error" gpurun_out/pytest_gpu.log | head -40
(for rep in 1 2; do
for v in r1 A B C D E F; do
NB_LIB_PATH=build/libnb_$v.so timeout 300 python tools/time_step.py 1048576:6 65536:60
done; done
) > gpurun_out/time_variants.txt 2>&1
cat gpurun_out/time_variants.txt
