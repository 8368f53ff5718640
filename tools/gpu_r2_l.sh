#!/bin/bash
# round 2, visit L: merge replay v3 (row-bucketed events, two-tier scans)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_collision_merge.py tests/test_gpu_dropin.py -m gpu -q > gpurun_out/pytest_merge.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_merge.log
grep -E "^(FAILED|ERROR)|passed|failed|exit|Error" gpurun_out/pytest_merge.log | head -20
timeout 200 python tests/tools/time_merge.py > gpurun_out/time_merge.txt 2>&1; cat gpurun_out/time_merge.txt
timeout 240 python tests/tools/time_merge.py big > gpurun_out/time_merge_big.txt 2>&1; tail -n 2 gpurun_out/time_merge_big.txt
