#!/bin/bash
# round 2, visit B: first run of the group-origin j-stream + speculative drift
mkdir -p gpurun_out
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -5 gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_fullsize.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -40 gpurun_out/pytest_gpu.log
timeout 600 python tests/tools/evolved_accuracy.py disk4096 disk3d4096 plummer16384 plummer65536 > gpurun_out/evolved_accuracy.txt 2>&1
cat gpurun_out/evolved_accuracy.txt
timeout 300 python bench.py --steps 3 --warmup 3 --no-tree --cpu-sample 0 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
