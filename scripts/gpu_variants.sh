#!/bin/bash
# dev: gpu tests with the default kernels, then the grid benches for each value of $VAR_NAME in $VARS
mkdir -p gpurun_out
VAR_NAME=${VAR_NAME:-B2FWN_FWN}
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
for v in ${VARS:-0 1 2}; do
for wl in ${WLS:-c1 c3_band c3_full c5_band}; do
  env $VAR_NAME=$v timeout 600 python bench.py --workload $wl --steps 3 --warmup 2 --no-e2e --no-cpu > gpurun_out/v${v}_$wl.json 2> gpurun_out/v${v}_$wl.err; rc=$?
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/v${v}_$wl.json").read().strip().splitlines()[-1]); print("$VAR_NAME=$v $wl", "%.1f Mvox/s"%(d["value"]/1e6), "%.2f ms"%d["ms_per_step"], "chk %.6f"%d["checksum"], d["clocks"]["sm_mhz"])
except Exception as e: print("$v $wl rc=$rc parse fail", e); print(open("gpurun_out/v${v}_$wl.err").read()[-1500:])
PY
done; done
