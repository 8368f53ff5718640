#!/bin/bash
# round 2, visit Z (2 GPUs): regrouping along the curve during a run (NB_RESORT_EVERY small), resume across it, both transports
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_dropin.py -m gpu -q -rs > gpurun_out/pytest_multi_z.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi_z.log
grep -v "^SKIPPED" gpurun_out/pytest_multi_z.log | tail -n 40
