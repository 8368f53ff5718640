#!/bin/bash
# round 2, visit R (1 GPU): small-world frame times before/after the deeper near-path unroll at R <= 2; headline unchanged?
mkdir -p gpurun_out
(NB_LIB_PATH=build/libnb_prev.so timeout 200 python tools/time_configs.py small | sed 's/^/prev /'
 timeout 200 python tools/time_configs.py small
 for rep in 1 2; do timeout 200 python tools/time_step.py 1048576:6 65536:60; done) > gpurun_out/time_small.txt 2>&1; cat gpurun_out/time_small.txt
