#!/bin/bash
# round 2, scale run (gpurun --gpus 8): bench at N = 8, 4, 2 (resident volumes under torchrun + the in-library multi-device e2e),
# the adapter on 8 devices, the C5-sized offset on 8 devices, the multi-device tests, ncu of the PEERS kernel at N = 2
O=gpurun_out/r2scale; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/gpu.txt; nvidia-smi topo -m >> $O/gpu.txt 2>&1; nproc >> $O/gpu.txt; free -g >> $O/gpu.txt
show() { python - <<PY
import json
try:
    d=json.loads(open("$1").read().strip().splitlines()[-1]); e=d.get("e2e") or {}
    print("$1", "%.2f Gvox/s"%(d["value"]/1e9), "%.1f ms"%d["ms_per_step"], "kernel/rank", [round(x,1) for x in (d.get("kernel_ms_per_rank") or [])], "| e2e %.1f ms = %.2f Gvox/s"%(e.get("ms_per_step",0), e.get("value",0)/1e9), "pinned", (e.get("pinned_buffer") or {}).get("ms_per_step"), "same", e.get("equals_resident_volume_on_sample_layers"))
except Exception as ex: print("parse fail $1", ex); print(open("$2").read()[-2500:])
PY
}
for n in 8 4 2; do
  EXTRA=""; [ $n != 8 ] && EXTRA="--no-pinned"
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 3 --warmup 3 $EXTRA > $O/scale_n$n.json 2> $O/scale_n$n.err; echo "n=$n rc=$?"; show $O/scale_n$n.json $O/scale_n$n.err
done
timeout 600 python scripts/adapter_bench.py c5_band 0,1,2,3,4,5,6,7 > $O/adapter_bench_c5_band_8dev.txt 2>&1; cat $O/adapter_bench_c5_band_8dev.txt
timeout 600 python scripts/diag_offset_devices.py > $O/diag_offset_devices.txt 2>&1; cat $O/diag_offset_devices.txt
timeout 600 python -m pytest tests -x -q -m gpu -k "several_devices or over_several or cpp_adapter" > $O/pytest_multi.log 2>&1; tail -3 $O/pytest_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 --no-python bash scripts/ncu_rank0.sh $O/ncu_peers_n2_rank0.csv --gpus 2 --steps 1 --warmup 1 --no-e2e --no-cpu > $O/ncu_peers_n2.log 2>&1; echo "ncu peers rc=$?"; grep -v "^==" $O/ncu_peers_n2_rank0.csv | awk -F'","' 'NR>1{print $5, $(NF-2), $NF}' | cut -c1-200 | tail -20
