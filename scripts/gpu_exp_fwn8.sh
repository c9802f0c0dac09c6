#!/bin/bash
# Experiment of DESIGN 10.1: kGridFwn8 (8 voxels per lane, B2FWN_FWN8=1) against kGridFwn5, both in plain linear brick order.
# 1. parity: the whole GPU suite with the switch on (per-voxel arithmetic is shared, results must be bit-identical);
# 2. timing: c3_band and c5_band, kernels resident, 3 steps each.
mkdir -p gpurun_out
B2FWN_FWN8=1 timeout 300 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/fwn8_pytest.log
for w in c3_band c5_band; do
  for v in 0 1; do
    B2FWN_LINEAR=1 B2FWN_FWN8=$v timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-e2e \
      > gpurun_out/fwn8_${w}_v$v.json 2> gpurun_out/fwn8_${w}_v$v.err
    python - <<PY
import json
d = json.loads(open("gpurun_out/fwn8_${w}_v$v.json").read().strip().splitlines()[-1])
print("$w FWN8=$v", d.get("kernel_ms_per_step"), "ms kernels,", d.get("ms_per_step"), "ms/step")
PY
  done
done
