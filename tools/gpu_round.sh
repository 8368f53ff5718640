#!/bin/bash
# One GPU-box visit: parity tests, bench (both arms), launch list and a full ncu capture of the force kernel.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
BCMD="python bench.py --steps 1 --warmup 3 --cpu-sample 0"
timeout 300 $BCMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $BCMD > gpurun_out/ncu_list.log 2>&1
timeout 300 $BCMD > gpurun_out/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:force_kernel -s 3 -c 1 -f -o gpurun_out/prof_force $BCMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json; cat gpurun_out/bench_ref.json
# the reference-solver (Barnes-Hut) pass: timing, launch list, full captures of its two kernels
TCMD="python tools/bh_pass.py 1048576 3 plummer"
timeout 300 $TCMD > gpurun_out/bh_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file gpurun_out/bh_launches.csv $TCMD > gpurun_out/bh_ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bh_walk -s 1 -c 1 -f -o gpurun_out/prof_bh_walk $TCMD > gpurun_out/bh_ncu_walk.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bh_build -s 1 -c 1 -f -o gpurun_out/prof_bh_build $TCMD > gpurun_out/bh_ncu_build.log 2>&1
cat gpurun_out/bh_plain.log
for n in 1000 4096 16384 65536 262144; do timeout 120 python tools/bh_pass.py $n 20 plummer; done > gpurun_out/bh_sizes.txt 2>&1; cat gpurun_out/bh_sizes.txt
