#!/bin/bash
# round 2, call N: lower tree levels in one block per subtree (bit-identical tree: the GPU suite), build times with and without
O=gpurun_out/r2n; mkdir -p $O
timeout 1800 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -8 $O/pytest_gpu.log
echo "subtree kernel:"; python scripts/diag_tree_build.py 2>&1 | tee $O/diag_tree_build_subtrees.txt
echo "global passes only:"; B2FWN_NO_SUBTREES=1 python scripts/diag_tree_build.py 2>&1 | tee $O/diag_tree_build_global.txt
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'kSubtrees' -c 4 --csv --log-file $O/launches_subtrees.csv python scripts/diag_tree_build.py > /dev/null 2>&1; grep kSubtrees $O/launches_subtrees.csv | awk -F'","' '{print $NF}'
