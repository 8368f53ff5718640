#!/bin/bash
# round 2, visit O (1 GPU): full GPU suite on the final library, evolved-state accuracy, planner experiments, bench, config 5
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|exit" gpurun_out/pytest_gpu.log | head -20
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -n 2 gpurun_out/smoke.log
timeout 600 python tests/tools/evolved_accuracy.py disk4096 disk3d4096 plummer16384 plummer65536 > gpurun_out/evolved_accuracy.txt 2>&1; cat gpurun_out/evolved_accuracy.txt
(for r in 4 2; do for s in 5 9 13 18 27; do NB_FORCE_R=$r NB_FORCE_NSEG=$s timeout 100 python tools/time_step.py 65536:80 | sed "s/^/R=$r nseg=$s /"; done; done
 timeout 100 python tools/time_step.py 65536:80 4096:200 1024:200 300:200) > gpurun_out/time_plan.txt 2>&1; cut -c1-150 gpurun_out/time_plan.txt
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_n1.json
timeout 900 python bench.py --config 5 --gpus 1 --steps 2 --warmup 3 --cpu-sample 0 > gpurun_out/config5_n1.json 2> gpurun_out/config5_n1.err; echo "config5 exit $?"; cut -c1-300 gpurun_out/config5_n1.json
