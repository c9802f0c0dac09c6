#!/bin/bash
# round 2, call K (2 GPUs): marching cubes by slabs / over devices, bench N=2 with the host-only barrier, MC kernel profile
O=gpurun_out/r2k; mkdir -p $O
timeout 1800 python -m pytest tests -x -q -m gpu --durations=5 > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -12 $O/pytest_gpu.log
show() { python - <<PY
import json
try:
    d=json.loads(open("$1").read().strip().splitlines()[-1]); e=d.get("e2e") or {}
    print("$1", "%.2f Gvox/s"%(d["value"]/1e9), "%.1f ms"%d["ms_per_step"], "kernel/rank", d.get("kernel_ms_per_rank"), "| e2e %.1f ms"%e.get("ms_per_step",0), "pinned", (e.get("pinned_buffer") or {}).get("ms_per_step"), "same", e.get("equals_resident_volume_on_sample_layers"), "mc_ms", d.get("mc_ms"), "sdf_ms", d.get("sdf_ms"))
except Exception as ex: print("parse fail $1", ex); print(open("$2").read()[-2500:])
PY
}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err; show $O/bench_n2.json $O/bench_n2.err
timeout 600 python bench.py --workload c5_offset --steps 2 --warmup 1 --no-cpu > $O/bench_c5_offset.json 2> $O/bench_c5_offset.err; show $O/bench_c5_offset.json $O/bench_c5_offset.err
python scripts/diag_offset_devices.py > $O/diag_offset_devices.txt 2>&1; cat $O/diag_offset_devices.txt
ncu --set full --clock-control none --import-source on -k regex:'kMcClassify|kMcEmit' -c 2 -f -o $O/prof_c3_offset_mc python bench.py --workload c3_offset --steps 1 --warmup 0 --no-cpu > $O/ncu_mc.log 2>&1; tail -1 $O/ncu_mc.log
