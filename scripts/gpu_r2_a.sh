#!/bin/bash
# round 2, first call: GPU tests with the v6 winding walk as default, then v6 vs v5 on the bench workloads
mkdir -p gpurun_out/r2
O=gpurun_out/r2
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > $O/gpu.txt 2>&1
timeout 900 python -m pytest tests -x -q -m gpu > $O/pytest_gpu_a.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu_a.log
tail -8 $O/pytest_gpu_a.log
for wl in c3_band c5_band c1 c3_full; do
  for v in 6 5; do
    if [ $v = 5 ]; then export B2FWN_FWN5=1; else unset B2FWN_FWN5; fi
    timeout 600 python bench.py --workload $wl --steps 3 --warmup 2 --no-e2e --no-cpu > $O/a_${wl}_v$v.json 2> $O/a_${wl}_v$v.err; echo "$wl v$v exit $?"
    python - <<PY
import json
try:
    d=json.loads(open("$O/a_${wl}_v$v.json").read().strip().splitlines()[-1]); print("$wl v$v", "%.1f Mvox/s"%(d["value"]/1e6), "%.2f ms"%d["ms_per_step"], "checksum", d["checksum"], d["clocks"])
except Exception as e: print("parse fail", e); print(open("$O/a_${wl}_v$v.err").read()[-2000:])
PY
  done
done
unset B2FWN_FWN5
