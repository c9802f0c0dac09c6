#!/bin/bash
# round 2, call M: lifecycle timing of the small workloads, default bench line, c2/c4 lines, c3_offset
O=gpurun_out/r2m; mkdir -p $O
python scripts/diag_lifecycle.py > $O/diag_lifecycle.txt 2>&1; cat $O/diag_lifecycle.txt
timeout 900 python bench.py --steps 3 --warmup 3 --cpu-budget 8 > $O/bench_c5_band.json 2> $O/bench_c5_band.err; python - <<PY
import json
try:
    d=json.loads(open("$O/bench_c5_band.json").read().strip().splitlines()[-1]); e=d["e2e"]; print("c5_band", "%.1f ms"%d["ms_per_step"], "e2e pageable %.1f ms"%e["ms_per_step"], "pinned %.1f"%e["pinned_buffer"]["ms_per_step"], "parity", d["parity"]["ok"], "frac", d["roofline"]["frac"])
except Exception as ex: print("parse fail", ex); print(open("$O/bench_c5_band.err").read()[-2000:])
PY
for wl in c2 c4 c3_offset; do
  timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --cpu-budget 8 > $O/bench_$wl.json 2> $O/bench_$wl.err; python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$wl.json").read().strip().splitlines()[-1]); print("$wl", "%.2f ms"%d["ms_per_step"], "e2e %.1f ms"%d["e2e"]["ms_per_step"], d.get("parity"), "frac", d["roofline"]["frac"])
except Exception as e: print("parse fail", e); print(open("$O/bench_$wl.err").read()[-3000:])
PY
done
