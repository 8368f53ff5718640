#!/bin/bash
# round 2, visit V (1 GPU): launch plan that prefers small R for worlds of <= 128 source tiles:
# whole GPU suite, config 3 + small worlds, step times across sizes, then the bench line.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|exit" gpurun_out/pytest_gpu.log | head
(timeout 200 python tools/time_configs.py c3 small
 timeout 200 python tools/time_step.py 1048576:6 262144:20 131072:30 65536:60 32768:80 16384:100 4096:200 1024:200) > gpurun_out/time_v.txt 2>&1; cat gpurun_out/time_v.txt
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; tail -c 1500 gpurun_out/bench_n1.json
