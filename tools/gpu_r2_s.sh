#!/bin/bash
# round 2, visit S (1 GPU): "mid" kernel variants (near path unrolled deeper for worlds of <= 256 source tiles)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/pytest_parity.log 2>&1; tail -n 2 gpurun_out/pytest_parity.log
(timeout 200 python tools/time_configs.py small
 for rep in 1 2; do timeout 200 python tools/time_step.py 1048576:6 131072:40 65536:60 16384:100; done
 timeout 200 python tools/time_configs.py c3) > gpurun_out/time_mid.txt 2>&1; cat gpurun_out/time_mid.txt
