#!/bin/bash
# round 2, visit W (1 GPU): allocation-free diagnostics entry point: its test alone in a fresh process, then the whole suite
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone exit $?" >> gpurun_out/pytest_noalloc_alone.log
tail -15 gpurun_out/pytest_noalloc_alone.log
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|exit" gpurun_out/pytest_gpu.log | head
