#!/bin/bash
# round 2, call E: leaf compaction in kGridFwn6 (results must be bit-identical: checksums), geometric host chunks, 8 copier threads
O=gpurun_out/r2e; mkdir -p $O
timeout 1800 python -m pytest tests -x -q -m gpu --durations=8 > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -14 $O/pytest_gpu.log
for wl in c5_band c3_band c1; do
  timeout 600 python bench.py --workload $wl --steps 3 --warmup 2 --no-cpu --no-pinned > $O/q_$wl.json 2> $O/q_$wl.err; python - <<PY
import json
try:
    d=json.loads(open("$O/q_$wl.json").read().strip().splitlines()[-1]); print("$wl", "%.1f ms"%d["ms_per_step"], "checksum", d["checksum"], "e2e", "%.1f ms"%d["e2e"]["ms_per_step"], d["e2e"]["checksum"], "call kernels", d["e2e"]["kernel_ms_of_the_call"])
except Exception as e: print("parse fail", e); print(open("$O/q_$wl.err").read()[-3000:])
PY
done
timeout 600 python scripts/adapter_bench.py c5_band 0 > $O/adapter_bench_c5_band.txt 2>&1; cat $O/adapter_bench_c5_band.txt
CMD="python bench.py --workload c5_band --steps 1 --warmup 1 --no-e2e --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:kGridFwn6 -c 1 -f -o $O/prof_c5_band_fwn6b $CMD > $O/ncu_full_c5_band.log 2>&1
echo "full exit $?"; tail -2 $O/ncu_full_c5_band.log
