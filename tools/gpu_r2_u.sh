#!/bin/bash
# round 2, visit U (1 GPU): targets per thread (R) against N for mid-size worlds
mkdir -p gpurun_out
(for r in 4 2 1; do NB_FORCE_R=$r timeout 200 python tools/time_step.py 16384:100 32768:100 65536:60 131072:40 262144:20 | sed "s/^/R=$r /"; done) > gpurun_out/time_r_scan.txt 2>&1; cut -c1-120 gpurun_out/time_r_scan.txt
