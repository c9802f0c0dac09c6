#!/bin/bash
# round 2, call J: distance-walk occupancy variants in full-range mode, e2e with the refined tail, MC kernel times
O=gpurun_out/r2j; mkdir -p $O
timeout 1800 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
for wl in c3_full c1 c3_band; do
  for mb in 8 6 5 4; do
    B2FWN_DIST_MINB=$mb timeout 600 python bench.py --workload $wl --steps 3 --warmup 2 --no-cpu --no-e2e > $O/q_${wl}_minb$mb.json 2> $O/q_${wl}_minb$mb.err; python - <<PY
import json
try:
    d=json.loads(open("$O/q_${wl}_minb$mb.json").read().strip().splitlines()[-1]); print("$wl minb=$mb", "%.2f ms"%d["ms_per_step"], "checksum", d["checksum"])
except Exception as e: print("parse fail", e); print(open("$O/q_${wl}_minb$mb.err").read()[-2000:])
PY
  done
done
timeout 900 python bench.py --steps 3 --warmup 2 --no-cpu > $O/bench_c5_band.json 2> $O/bench_c5_band.err; python - <<PY
import json
try:
    d=json.loads(open("$O/bench_c5_band.json").read().strip().splitlines()[-1]); e=d["e2e"]; print("c5_band", "%.1f ms"%d["ms_per_step"], "e2e pageable %.1f ms"%e["ms_per_step"], "pinned %.1f"%e["pinned_buffer"]["ms_per_step"], "call kernels %.1f"%e["kernel_ms_of_the_call"], e["equals_resident_volume_on_sample_layers"])
except Exception as ex: print("parse fail", ex); print(open("$O/bench_c5_band.err").read()[-2000:])
PY
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'kMc|DeviceScan' -c 40 --csv --log-file $O/launches_mc_c3_offset.csv python bench.py --workload c3_offset --steps 1 --warmup 1 --no-cpu > $O/ncu_launches_mc.log 2>&1
python - <<PY
import csv
rows=list(csv.reader(open("$O/launches_mc_c3_offset.csv")))
hi=[i for i,r in enumerate(rows) if 'Kernel Name' in r][0]; h=rows[hi]; kn=h.index('Kernel Name'); mv=h.index('Metric Value')
for r in rows[hi+1:]:
    if len(r)>mv: print(r[kn][:60], float(r[mv].replace(',',''))/1e3, "us")
PY
