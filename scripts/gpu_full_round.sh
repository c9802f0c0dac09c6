#!/bin/bash
# one gpurun call: gpu tests, smoke, default bench (+ reference arm), diagnostic benches, ncu launch list + full captures
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc > gpurun_out/nproc.txt; free -g >> gpurun_out/nproc.txt
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "default exit $?"; cut -c1-600 gpurun_out/bench_default.json; tail -3 gpurun_out/bench_default.err
timeout 600 python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference exit $?"; cut -c1-300 gpurun_out/bench_reference.json
for wl in c1 c3_band c3_full; do
  timeout 600 python bench.py --workload $wl --steps 3 --warmup 3 --cpu-budget 5 > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "$wl exit $?"; cut -c1-400 gpurun_out/bench_$wl.json; tail -3 gpurun_out/bench_$wl.err
done
WL=c3_band
CMD="python bench.py --workload $WL --steps 2 --warmup 1 --no-e2e --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_$WL.csv $CMD > gpurun_out/ncu_launches_$WL.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'kGridDist|kGridFwn5' -c 2 -f -o gpurun_out/prof_$WL $CMD > gpurun_out/ncu_full_$WL.log 2>&1
tail -3 gpurun_out/ncu_full_$WL.log
# dominant kernel at the bench workload: full capture (DRAM traffic for roofline.traffic)
ncu --set full --clock-control none --import-source on -k regex:kGridFwn5 -c 1 -f -o gpurun_out/prof_c5_band_fwn5 python bench.py --workload c5_band --steps 1 --warmup 0 --no-e2e --no-cpu > gpurun_out/ncu_full_c5_band.log 2>&1
tail -2 gpurun_out/ncu_full_c5_band.log
python scripts/diag_queries.py > gpurun_out/diag_queries.txt 2>&1; tail -2 gpurun_out/diag_queries.txt
python scripts/diag_project.py > gpurun_out/diag_project.txt 2>&1; tail -2 gpurun_out/diag_project.txt
python scripts/diag_e2e.py c5_band > gpurun_out/diag_e2e_c5.txt 2>&1; tail -3 gpurun_out/diag_e2e_c5.txt
ncu --set full --clock-control none --import-source on -k regex:kFindProjections -c 1 -f -o gpurun_out/prof_project python scripts/diag_project.py > gpurun_out/ncu_full_project.log 2>&1
tail -2 gpurun_out/ncu_full_project.log
ls -la gpurun_out
