#!/bin/bash
# tests + short benches of the grid kernel (outputs under gpurun_out/)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; tail -12 gpurun_out/pytest_gpu.log
for wl in ${WLS:-c1 c3_band c3_full}; do
  timeout 600 python bench.py --workload $wl --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/q_$wl.json 2> gpurun_out/q_$wl.err; echo "$wl exit $?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/q_$wl.json").read().strip().splitlines()[-1]); print("$wl", "%.1f Mvox/s"%(d["value"]/1e6), "%.2f ms"%d["ms_per_step"], d["clocks"])
except Exception as e: print("parse fail", e); print(open("gpurun_out/q_$wl.err").read()[-2000:])
PY
done
