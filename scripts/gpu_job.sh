#!/bin/bash
mkdir -p gpurun_out
for w in c3_band c5_band; do B2FWN_DEBUG=1 python bench.py --workload $w --steps 1 --warmup 0 --no-e2e --no-cpu 2>&1 | grep "b2fwn order"; done
VAR_NAME=B2FWN_LINEAR VARS="0 1" WLS="c1 c3_band c5_band" bash scripts/gpu_variants.sh
for wl in c3_band; do
ncu --metrics gpu__time_duration.sum,sm__cycles_active.avg,sm__cycles_elapsed.avg,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:'kGrid|kBrick|kGroup' -c 12 --csv --log-file gpurun_out/${wl}_kernels.csv python bench.py --workload $wl --steps 1 --warmup 0 --no-e2e --no-cpu > gpurun_out/ncu_$wl.log 2>&1
python - <<PY
import csv
rows=list(csv.reader(l for l in open('gpurun_out/${wl}_kernels.csv') if l.startswith('"')))
h=rows[0]; cur=None; out=[]
for r in rows[1:]:
    k=r[h.index('ID')]
    if k!=cur: cur=k; out.append([r[h.index('Kernel Name')][:34]])
    out[-1].append(r[h.index('Metric Value')])
for o in out: print('$wl',*o)
PY
done
