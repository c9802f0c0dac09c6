#!/bin/bash
# round 2, call G (2 GPUs): per-device parallel create, MC fixes, half-edge tables
O=gpurun_out/r2g; mkdir -p $O
timeout 1800 python -m pytest tests -x -q -m gpu --durations=6 > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -12 $O/pytest_gpu.log
show() { python - <<PY
import json
try:
    d=json.loads(open("$1").read().strip().splitlines()[-1]); e=d.get("e2e") or {}
    print("$1", "%.2f Gvox/s"%(d["value"]/1e9), "%.1f ms"%d["ms_per_step"], "kernel/rank", d.get("kernel_ms_per_rank"), "| e2e %.1f ms"%e.get("ms_per_step",0), "pinned", (e.get("pinned_buffer") or {}).get("ms_per_step"), "same", e.get("equals_resident_volume_on_sample_layers"), "| parity", (d.get("parity") or {}), "mc_ms", d.get("mc_ms"), "sdf_ms", d.get("sdf_ms"), d.get("roofline",{}).get("frac"))
except Exception as ex: print("parse fail $1", ex); print(open("$2").read()[-2500:])
PY
}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err; show $O/bench_n2.json $O/bench_n2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 --gather nccl --no-e2e > $O/bench_n2_nccl.json 2> $O/bench_n2_nccl.err; show $O/bench_n2_nccl.json $O/bench_n2_nccl.err
timeout 600 python scripts/adapter_bench.py c5_band 0,1 > $O/adapter_bench_c5_band_2dev.txt 2>&1; cat $O/adapter_bench_c5_band_2dev.txt
timeout 600 python bench.py --workload c3_offset --steps 3 --warmup 2 > $O/bench_c3_offset.json 2> $O/bench_c3_offset.err; show $O/bench_c3_offset.json $O/bench_c3_offset.err
timeout 600 python bench.py --workload c5_offset --steps 2 --warmup 1 --no-cpu > $O/bench_c5_offset.json 2> $O/bench_c5_offset.err; show $O/bench_c5_offset.json $O/bench_c5_offset.err
python scripts/diag_topology.py > $O/diag_topology.txt 2>&1; cat $O/diag_topology.txt
