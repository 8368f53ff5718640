#!/bin/bash
# round 2, final single-GPU visit: whole GPU suite, then the driver's bench command twice (run-to-run spread)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|exit" gpurun_out/pytest_gpu.log | head
for k in a b; do
  timeout 600 python bench.py > gpurun_out/bench_n1_$k.json 2> gpurun_out/bench_n1_$k.err; echo "bench $k exit $?"
  python - <<PY
import json
d = json.load(open("gpurun_out/bench_n1_$k.json"))
print({q: d.get(q) for q in ("ms_per_step", "value", "gpu_launches")}, d["e2e"]["ms_per_step"], d["roofline"]["frac"], d["roofline"]["force_ms_per_step"], d["parity"]["ok"], d["config3"], d["general_mass"]["ms_per_step"], d["planar_2d"]["ms_per_step"], d["diagnostics"]["seconds"], d["clocks"])
PY
done
