#!/bin/bash
# multi-GPU evidence of round 3 (gpurun --gpus N): the multi-device tests by name, then bench lines at the given rank counts
# usage: scripts/gpu_r3_scale.sh <outdir> <n> [<n> ...]
OUT=gpurun_out/$1; shift
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_pipeline.py -v -m gpu -k "several_devices" > $OUT/pytest_multi_device.log 2>&1; echo "multi-device tests rc=$?"; tail -5 $OUT/pytest_multi_device.log
for n in "$@"; do
  if [ $n = 1 ]; then CMD="python bench.py"; else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n"; fi
  timeout 900 $CMD --steps 10 --warmup 3 --no-e2e --no-cpu > $OUT/scale_n$n.json 2> $OUT/scale_n$n.err; echo "n=$n rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("$OUT/scale_n$n.json").read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ["value","n_gpus","ms_per_step","kernel_ms_per_step","kernel_ms_per_rank","gather_exposed_ms_per_step","checksum"]}, d["roofline"].get("frac"))
except Exception as e: print("fail", e); print(open("$OUT/scale_n$n.err").read()[-1500:])
PY
done
