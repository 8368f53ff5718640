#!/bin/bash
# round 2, visit K: tests, step/merge timings, ncu captures summarised on the box (reports are too large to bring back)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_fullsize.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|exit|Error" gpurun_out/pytest_gpu.log | head -40
timeout 600 python tests/tools/time_merge.py big > gpurun_out/time_merge.txt 2>&1; cat gpurun_out/time_merge.txt
(for rep in 1 2; do timeout 300 python tools/time_step.py 1048576:6 65536:60 16384:100 4096:200 1024:200; done
 timeout 300 python tools/time_step.py 1048576:4 65536:60 2) > gpurun_out/time_step.txt 2>&1; cat gpurun_out/time_step.txt
BCMD="python bench.py --steps 2 --warmup 3 --cpu-sample 0 --no-sides --no-tree --no-parity"
timeout 300 $BCMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $BCMD > gpurun_out/ncu_list.log 2>&1
timeout 300 $BCMD > gpurun_out/plain2.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:force_kernel -s 4 -c 1 -f -o /tmp/prof_force $BCMD > gpurun_out/ncu_full.log 2>&1
ncu -i /tmp/prof_force.ncu-rep --page raw --csv > gpurun_out/ncu_force_raw.csv 2>/dev/null
ncu -i /tmp/prof_force.ncu-rep --page source --csv > /tmp/src.csv 2>/dev/null; head -c 3000000 /tmp/src.csv > gpurun_out/ncu_force_source_head.csv
TCMD="python tools/ncu_targets.py"
timeout 300 $TCMD > gpurun_out/plain3.log 2>&1 && \
timeout 1800 ncu --set full --clock-control none -k regex:force_kernel -s 1 -c 4 -f -o /tmp/prof_pot $TCMD > gpurun_out/ncu_pot.log 2>&1
ncu -i /tmp/prof_pot.ncu-rep --page raw --csv > gpurun_out/ncu_pot_raw.csv 2>/dev/null
ls -la /tmp/*.ncu-rep gpurun_out/; du -sh gpurun_out
