#!/bin/bash
# round-3 GPU jobs, one script with stages: scripts/gpu_r3.sh <outdir> <stage> [<stage> ...]
#   tests      pytest -m gpu (verbose names)              -> pytest_gpu.log
#   bench      default bench line (c5_band, N = 1)        -> bench_c5_band.json
#   launches   ncu launch list of the default bench cmd   -> ncu_launches_c5_band.csv
#   fwn6       full capture, ALL passes, of kGridFwn6 on the whole c5_band grid (application replay) -> ncu_full_kGridFwn6_c5_band.*
#   c3         full capture of both grid kernels at c3_band (kernel replay)                          -> ncu_full_c3_band.*
#   c3full     full capture of kGridDist at c3_full                                                  -> ncu_full_kGridDist_c3_full.*
OUT=gpurun_out/$1; shift
mkdir -p $OUT
for stage in "$@"; do
case $stage in
tests)
  timeout 1500 python -m pytest tests -x -q -m gpu -v -rs > $OUT/pytest_gpu.log 2>&1; echo "tests rc=$?"; tail -4 $OUT/pytest_gpu.log;;
bench)
  timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/bench_c5_band.json 2> $OUT/bench_c5_band.err; echo "bench rc=$?";;
launches)
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/ncu_launches_c5_band.csv \
    python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > $OUT/ncu_launches_c5_band.log 2>&1; echo "launches rc=$?";;
fwn6)
  python scripts/ncu_one_launch.py c5_band > $OUT/one_launch_plain.log 2>&1 && \
  timeout 2400 ncu --set full --clock-control none --import-source on --replay-mode application -k regex:kGridFwn6 -c 1 -f \
    -o $OUT/ncu_full_kGridFwn6_c5_band python scripts/ncu_one_launch.py c5_band > $OUT/ncu_full_kGridFwn6_c5_band.log 2>&1
  echo "fwn6 rc=$?"; tail -3 $OUT/ncu_full_kGridFwn6_c5_band.log;;
c3)
  python scripts/ncu_one_launch.py c3_band 2 > $OUT/one_launch_c3.log 2>&1 && \
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:kGrid -s 2 -c 2 -f -o $OUT/ncu_full_c3_band \
    python scripts/ncu_one_launch.py c3_band 2 > $OUT/ncu_full_c3_band.log 2>&1; echo "c3 rc=$?";;
c3full)
  python scripts/ncu_one_launch.py c3_full 1 > $OUT/one_launch_c3full.log 2>&1 && \
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:kGridDist -c 1 -f -o $OUT/ncu_full_kGridDist_c3_full \
    python scripts/ncu_one_launch.py c3_full 1 > $OUT/ncu_full_kGridDist_c3_full.log 2>&1; echo "c3full rc=$?";;
esac
done
for r in $OUT/*.ncu-rep; do [ -f "$r" ] || continue
  python scripts/ncu_summary.py $r > ${r%.ncu-rep}_summary.txt 2>&1
  ncu -i $r --page raw --csv > ${r%.ncu-rep}.csv 2>/dev/null
  python scripts/ncu_ops.py $r > ${r%.ncu-rep}_sass_ops.txt 2>&1
done
ls -la $OUT
