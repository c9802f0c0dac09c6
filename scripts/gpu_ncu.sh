#!/bin/bash
# launch list + one full capture of the top kernel (bench workload given as $1, default c3_band)
WL=${1:-c3_band}
mkdir -p gpurun_out
CMD="python bench.py --workload $WL --steps 2 --warmup 1 --no-e2e --no-cpu"
$CMD > gpurun_out/ncu_plain_$WL.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_$WL.csv $CMD > gpurun_out/ncu_launches_$WL.log 2>&1
$CMD > gpurun_out/ncu_plain2_$WL.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:kGrid -c 1 -f -o gpurun_out/prof_$WL $CMD > gpurun_out/ncu_full_$WL.log 2>&1
tail -3 gpurun_out/ncu_full_$WL.log
