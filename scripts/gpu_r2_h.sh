#!/bin/bash
# round 2, call H: Morton / leaf ordered queries (c2, c4 bench lines), adapter offsetMesh, MC launch list
O=gpurun_out/r2h; mkdir -p $O
timeout 1800 python -m pytest tests -x -q -m gpu --durations=5 > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -10 $O/pytest_gpu.log
oracle/_ref/adapter_check 2>&1 | grep -i "offsetMesh\|FAIL"
for wl in c2 c4; do
  timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --cpu-budget 10 > $O/bench_$wl.json 2> $O/bench_$wl.err; python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$wl.json").read().strip().splitlines()[-1]); print("$wl", "%.1f Mq/s"%(d["value"]/1e6), "%.2f ms"%d["ms_per_step"], "e2e %.1f ms"%d["e2e"]["ms_per_step"], "cpu %.2f Mq/s"%(d["cpu_baseline"]["value"]/1e6), d["parity"], "frac", d["roofline"]["frac"], d["roofline"].get("ops_per_query"))
except Exception as e: print("parse fail", e); print(open("$O/bench_$wl.err").read()[-3000:])
PY
done
B2FWN_NO_QUERY_SORT=1 timeout 300 python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu > $O/bench_c2_nosort.json 2> $O/bench_c2_nosort.err; python -c "
import json; d=json.loads(open('$O/bench_c2_nosort.json').read().strip().splitlines()[-1]); print('c2 input order: %.2f ms'%d['ms_per_step'])"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3_offset.csv python bench.py --workload c3_offset --steps 1 --warmup 1 --no-cpu > $O/ncu_launches_c3_offset.log 2>&1
grep -i "kMc\|DeviceScan" $O/launches_c3_offset.csv | awk -F'","' '{print $5, $NF}' | tail -12
