#!/bin/bash
# round 2, visit J: the bench as the driver runs it (1 GPU), its launch list, full ncu captures
mkdir -p gpurun_out
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"
cat gpurun_out/bench_n1.json; tail -n 5 gpurun_out/bench_n1.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
BCMD="python bench.py --steps 2 --warmup 3 --cpu-sample 0 --no-sides --no-tree --no-parity"
timeout 300 $BCMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $BCMD > gpurun_out/ncu_list.log 2>&1
timeout 300 $BCMD > gpurun_out/plain2.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:force_kernel -s 4 -c 1 -f -o gpurun_out/prof_force $BCMD > gpurun_out/ncu_full.log 2>&1
TCMD="python tools/ncu_targets.py"
timeout 300 $TCMD > gpurun_out/plain3.log 2>&1 && \
timeout 1800 ncu --set full --clock-control none --import-source on -k regex:force_kernel -s 1 -c 4 -f -o gpurun_out/prof_pot $TCMD > gpurun_out/ncu_pot.log 2>&1
cat gpurun_out/plain3.log; ls -la gpurun_out/*.ncu-rep; tail -n 3 gpurun_out/ncu_full.log gpurun_out/ncu_pot.log
