#!/bin/bash
# one gpurun call: tests, smoke, bench at several sizes (outputs under gpurun_out/)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc > gpurun_out/nproc.txt; free -g >> gpurun_out/nproc.txt
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
for wl in c1 c3_band c3_full; do
  timeout 600 python bench.py --workload $wl --steps 2 --warmup 1 --cpu-budget 5 > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "$wl exit $?"; cat gpurun_out/bench_$wl.json; tail -3 gpurun_out/bench_$wl.err
done
timeout 900 python bench.py --workload c5_band --steps 1 --warmup 1 --no-cpu > gpurun_out/bench_c5_band.json 2> gpurun_out/bench_c5_band.err; echo "c5_band exit $?"; cat gpurun_out/bench_c5_band.json; tail -3 gpurun_out/bench_c5_band.err
