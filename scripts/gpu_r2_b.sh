#!/bin/bash
# round 2: per-kernel launch lists (v6 and v5) and full captures of the v6 winding walk
O=gpurun_out/r2; mkdir -p $O
for wl in c3_band c5_band; do
  for v in 6 5; do
    if [ $v = 5 ]; then export B2FWN_FWN5=1; else unset B2FWN_FWN5; fi
    CMD="python bench.py --workload $wl --steps 1 --warmup 1 --no-e2e --no-cpu"
    ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_${wl}_v$v.csv $CMD > $O/ncu_launches_${wl}_v$v.log 2>&1
    echo "launches $wl v$v exit $?"
  done
done
unset B2FWN_FWN5
for wl in c5_band c3_band; do
  CMD="python bench.py --workload $wl --steps 1 --warmup 0 --no-e2e --no-cpu"
  ncu --set full --clock-control none --import-source on -k regex:kGridFwn6 -c 1 -f -o $O/prof_${wl}_fwn6 $CMD > $O/ncu_full_${wl}_fwn6.log 2>&1
  echo "full $wl exit $?"; tail -2 $O/ncu_full_${wl}_fwn6.log
done
ls -la $O
