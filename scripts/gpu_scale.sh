#!/bin/bash
# gpu tests + the scaling runs (gpurun --gpus 8): bench at N = 8, 4, 2, 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
for n in ${NS:-8 4 2 1}; do
  if [ $n = 1 ]; then CMD="python bench.py"; else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n"; fi
  timeout 900 $CMD --steps 3 --warmup 3 $EXTRA > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err; echo "n=$n rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/scale_n$n.json").read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ["value","n_gpus","ms_per_step","kernel_ms_per_step","gather_exposed_ms_per_step","wall_ms_per_step","gpu_launches"]}, "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"])
except Exception as e: print("fail", e); print(open("gpurun_out/scale_n$n.err").read()[-1500:])
PY
done
