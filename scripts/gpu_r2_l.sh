#!/bin/bash
# round 2, call L: warp-cooperative query kernels (c2, c4), cached mesher workspace (offset workloads, slabs)
O=gpurun_out/r2l; mkdir -p $O
timeout 1800 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -8 $O/pytest_gpu.log
for wl in c2 c4; do
  timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --cpu-budget 10 > $O/bench_$wl.json 2> $O/bench_$wl.err; python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$wl.json").read().strip().splitlines()[-1]); print("$wl", "%.1f Mq/s"%(d["value"]/1e6), "%.2f ms"%d["ms_per_step"], "e2e %.1f ms"%d["e2e"]["ms_per_step"], "cpu %.2f Mq/s"%(d["cpu_baseline"]["value"]/1e6), d["parity"], "frac", d["roofline"]["frac"])
except Exception as e: print("parse fail", e); print(open("$O/bench_$wl.err").read()[-3000:])
PY
  B2FWN_QUERY_THREADS=1 timeout 300 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu > $O/bench_${wl}_threads.json 2> $O/bench_${wl}_threads.err; python -c "
import json; d=json.loads(open('$O/bench_${wl}_threads.json').read().strip().splitlines()[-1]); print('$wl one thread per query: %.2f ms'%d['ms_per_step'])"
done
for wl in c3_offset c5_offset; do
timeout 600 python bench.py --workload $wl --steps 3 --warmup 2 --no-cpu > $O/bench_$wl.json 2> $O/bench_$wl.err; python -c "
import json; d=json.loads(open('$O/bench_$wl.json').read().strip().splitlines()[-1]); print('$wl sdf %.1f mc %.1f e2e %.1f'%(d['sdf_ms'], d['mc_ms'], d['e2e']['ms_per_step']), d['n_verts'], d['n_tris'], 'mc hbm frac', d['roofline']['frac'])"
done
python scripts/diag_offset_devices.py > $O/diag_offset_devices.txt 2>&1; cat $O/diag_offset_devices.txt
