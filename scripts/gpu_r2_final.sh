#!/bin/bash
# round 2, evidence run on one GPU: GPU suite, smoke, the default bench line + reference arm, the other workloads, launch list and
# full capture of the winding walk, tree build times.  Output: gpurun_out/r2final (copied to profiles/r02 afterwards)
O=gpurun_out/r2final; mkdir -p $O
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > $O/gpu.txt 2>&1; nproc >> $O/gpu.txt; free -g >> $O/gpu.txt
timeout 1800 python -m pytest tests -x -q -m gpu --durations=5 > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -10 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke exit $?" >> $O/smoke.log; tail -2 $O/smoke.log
python scripts/diag_tree_build.py > $O/diag_tree_build.txt 2>&1; cat $O/diag_tree_build.txt
timeout 900 python bench.py > $O/bench_c5_band.json 2> $O/bench_c5_band.err; echo "default exit $?"; cut -c1-400 $O/bench_c5_band.json; tail -3 $O/bench_c5_band.err
timeout 600 python bench.py --impl reference > $O/bench_reference_arm_c5_band.json 2> $O/bench_reference.err; echo "reference exit $?"; cut -c1-300 $O/bench_reference_arm_c5_band.json
for wl in c3_band c3_full c1 c2 c4 c3_offset; do
  timeout 600 python bench.py --workload $wl --steps 3 --warmup 3 --cpu-budget 6 > $O/bench_$wl.json 2> $O/bench_$wl.err; echo "$wl exit $?"; cut -c1-250 $O/bench_$wl.json; tail -2 $O/bench_$wl.err
done
timeout 600 python bench.py --workload c5_offset --steps 2 --warmup 1 --no-cpu > $O/bench_c5_offset.json 2> $O/bench_c5_offset.err; echo "c5_offset exit $?"; cut -c1-250 $O/bench_c5_offset.json
CMD="python bench.py --workload c5_band --steps 1 --warmup 1 --no-e2e --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/ncu_launches_c5_band.csv $CMD > $O/ncu_launches.log 2>&1; echo "launch list exit $?"
ncu --set full --clock-control none --import-source on -k regex:'kGridFwn6|kGridDist' -c 2 -f -o $O/prof_c5_band python bench.py --workload c5_band --steps 1 --warmup 0 --no-e2e --no-cpu > $O/ncu_full.log 2>&1; echo "full exit $?"; tail -1 $O/ncu_full.log
ls -la $O | tail -30
