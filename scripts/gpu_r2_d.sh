#!/bin/bash
# round 2, call D: the new host pipeline -- GPU suite (incl. pipeline + scale tests), bench with the pageable e2e leg, the adapter
# bench, staging ring vs cudaHostRegister
O=gpurun_out/r2d; mkdir -p $O
timeout 1800 python -m pytest tests -x -q -m gpu --durations=12 > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -25 $O/pytest_gpu.log
timeout 900 python bench.py > $O/bench_c5_band.json 2> $O/bench_c5_band.err; echo "default exit $?"; python - <<PY
import json
try:
    d=json.loads(open("$O/bench_c5_band.json").read().strip().splitlines()[-1]); print(d["value"]/1e9, d["ms_per_step"], "e2e", d["e2e"])
except Exception as e: print("parse fail", e); print(open("$O/bench_c5_band.err").read()[-3000:])
PY
timeout 600 python scripts/adapter_bench.py c5_band 0 byparts > $O/adapter_bench_c5_band.txt 2>&1; echo "adapter exit $?"; cat $O/adapter_bench_c5_band.txt
timeout 300 python scripts/adapter_bench.py c3_band 0 byparts > $O/adapter_bench_c3_band.txt 2>&1; cat $O/adapter_bench_c3_band.txt
B2FWN_HOST_OUT=register timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu --no-pinned > $O/bench_c5_band_register.json 2> $O/bench_c5_band_register.err; python - <<PY
import json
try:
    d=json.loads(open("$O/bench_c5_band_register.json").read().strip().splitlines()[-1]); print("register:", d["e2e"]["ms_per_step"])
except Exception as e: print("parse fail", e); print(open("$O/bench_c5_band_register.err").read()[-2000:])
PY
for t in 2 8; do
B2FWN_COPY_THREADS=$t timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu --no-pinned > $O/bench_c5_band_t$t.json 2> $O/bench_c5_band_t$t.err; python - <<PY
import json
try:
    d=json.loads(open("$O/bench_c5_band_t$t.json").read().strip().splitlines()[-1]); print("copy threads $t:", d["e2e"]["ms_per_step"])
except Exception as e: print("parse fail", e); print(open("$O/bench_c5_band_t$t.err").read()[-2000:])
PY
done
ls -la $O
