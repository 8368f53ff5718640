#!/bin/bash
# round 2, visit T (1 GPU): final library: whole GPU suite, config-3 step + diagnostics by forced R, headline check
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|exit" gpurun_out/pytest_gpu.log | head
(for r in 0 4 3 2 1; do NB_FORCE_R=$r timeout 100 python tools/time_configs.py c3 | sed "s/^/R=$r /"; done
 timeout 200 python tools/time_step.py 1048576:6 65536:60 16384:100) > gpurun_out/time_c3.txt 2>&1; cat gpurun_out/time_c3.txt
