#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --cpu-sample 0 > gpurun_out/bench_eqm.json 2> gpurun_out/bench.err
NB_FORCE_GENERIC=1 timeout 600 python bench.py --steps 5 --warmup 3 --cpu-sample 0 > gpurun_out/bench_generic.json 2>> gpurun_out/bench.err
tail -4 gpurun_out/pytest_gpu.log; python - <<'PY'
import json
for f in ("bench_eqm","bench_generic"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json")); print(f, d["value"], d["pct_fp32_peak"], d["ms_per_step"], d["e2e"]["value"], d["clocks"])
    except Exception as e: print(f, "failed", e)
PY
tail -5 gpurun_out/bench.err
