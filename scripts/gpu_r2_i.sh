#!/bin/bash
# round 2, call I: windowed uniform loop of kGridFwn6 against the node-by-node loop (bit-identical checksums), MC classify fast path
O=gpurun_out/r2i; mkdir -p $O
timeout 1800 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
for wl in c5_band c3_band c1 c3_full; do
  for nw in 0 1; do
    B2FWN_NOWIN=$nw timeout 600 python bench.py --workload $wl --steps 3 --warmup 2 --no-cpu --no-e2e > $O/q_${wl}_nowin$nw.json 2> $O/q_${wl}_nowin$nw.err; python - <<PY
import json
try:
    d=json.loads(open("$O/q_${wl}_nowin$nw.json").read().strip().splitlines()[-1]); print("$wl nowin=$nw", "%.2f ms"%d["ms_per_step"], "checksum", d["checksum"])
except Exception as e: print("parse fail", e); print(open("$O/q_${wl}_nowin$nw.err").read()[-3000:])
PY
  done
done
timeout 600 python bench.py --workload c3_offset --steps 3 --warmup 2 --no-cpu > $O/bench_c3_offset.json 2> $O/bench_c3_offset.err; python -c "
import json; d=json.loads(open('$O/bench_c3_offset.json').read().strip().splitlines()[-1]); print('c3_offset sdf', d['sdf_ms'], 'mc', d['mc_ms'], d['n_verts'], d['n_tris'])"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3_offset.csv python bench.py --workload c3_offset --steps 1 --warmup 1 --no-cpu > $O/ncu_launches_c3_offset.log 2>&1
python - <<PY
import csv
rows=list(csv.reader(open("$O/launches_c3_offset.csv")))
hi=[i for i,r in enumerate(rows) if 'Kernel Name' in r][0]; h=rows[hi]; kn=h.index('Kernel Name'); mv=h.index('Metric Value')
for r in rows[hi+1:]:
    if len(r)>mv and ('kMc' in r[kn] or 'Scan' in r[kn]): print(r[kn][:60], r[mv])
PY
ncu --set full --clock-control none --import-source on -k regex:kGridFwn6 -c 1 -f -o $O/prof_c5_band_fwn6w python bench.py --workload c5_band --steps 1 --warmup 0 --no-e2e --no-cpu > $O/ncu_full_c5_band.log 2>&1; tail -1 $O/ncu_full_c5_band.log
