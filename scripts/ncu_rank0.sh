#!/bin/bash
# torchrun --no-python wrapper: rank 0 runs bench.py under ncu (the PEERS variant of the winding walk: duration, L2 write sectors,
# NVLink bytes), the other ranks run it plainly.  usage: torchrun ... --no-python scripts/ncu_rank0.sh OUT.csv <bench.py args>
OUT=$1; shift
if [ "$LOCAL_RANK" = "0" ]; then
  exec ncu --metrics gpu__time_duration.sum,lts__t_sectors_op_write.sum,lts__t_sectors_op_read.sum,dram__bytes_read.sum,dram__bytes_write.sum,nvltx__bytes.sum,nvlrx__bytes.sum,smsp__inst_executed.sum \
       --clock-control none -k regex:kGridFwn6 -c 2 --csv --log-file "$OUT" python bench.py "$@"
else
  exec python bench.py "$@"
fi
