#!/bin/bash
# round 2, call C: GPU suite, the default bench line with the parity block, c5_full (default options at the target size),
# launch list + full capture of the v6 winding walk at c5_band
O=gpurun_out/r2c; mkdir -p $O
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > $O/gpu.txt 2>&1
nproc > $O/nproc.txt; free -g >> $O/nproc.txt
timeout 1200 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -6 $O/pytest_gpu.log
timeout 900 python bench.py > $O/bench_c5_band.json 2> $O/bench_c5_band.err; echo "default exit $?"; cut -c1-700 $O/bench_c5_band.json; tail -3 $O/bench_c5_band.err
timeout 1500 python bench.py --workload c5_full --steps 1 --warmup 1 --cpu-budget 12 > $O/bench_c5_full.json 2> $O/bench_c5_full.err; echo "c5_full exit $?"; cut -c1-700 $O/bench_c5_full.json; tail -3 $O/bench_c5_full.err
for wl in c3_band c3_full c1; do
  timeout 600 python bench.py --workload $wl --steps 3 --warmup 3 --cpu-budget 6 > $O/bench_$wl.json 2> $O/bench_$wl.err; echo "$wl exit $?"; cut -c1-300 $O/bench_$wl.json; tail -3 $O/bench_$wl.err
done
CMD="python bench.py --workload c5_band --steps 1 --warmup 1 --no-e2e --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_c5_band.csv $CMD > $O/ncu_launches_c5_band.log 2>&1
echo "launches exit $?"
ncu --set full --clock-control none --import-source on -k regex:kGridFwn6 -c 1 -f -o $O/prof_c5_band_fwn6 python bench.py --workload c5_band --steps 1 --warmup 0 --no-e2e --no-cpu > $O/ncu_full_c5_band.log 2>&1
echo "full exit $?"; tail -2 $O/ncu_full_c5_band.log
ls -la $O
