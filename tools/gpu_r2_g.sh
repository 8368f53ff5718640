#!/bin/bash
# round 2, visit G: ncu --set full of the force kernel, round-1 build vs current, same command
mkdir -p gpurun_out
CMD="python tools/time_step.py 1048576:1"
NB_LIB_PATH=build/libnb_r1.so timeout 200 $CMD > gpurun_out/plain_r1.log 2>&1 && \
NB_LIB_PATH=build/libnb_r1.so timeout 900 ncu --set full --clock-control none --import-source on -k regex:force_kernel -s 2 -c 1 -f -o gpurun_out/prof_r1 $CMD > gpurun_out/ncu_r1.log 2>&1
timeout 200 $CMD > gpurun_out/plain_new.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:force_kernel -s 2 -c 1 -f -o gpurun_out/prof_new $CMD > gpurun_out/ncu_new.log 2>&1
tail -3 gpurun_out/ncu_r1.log gpurun_out/ncu_new.log; ls -la gpurun_out/*.ncu-rep
