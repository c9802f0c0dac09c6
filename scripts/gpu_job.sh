#!/bin/bash
mkdir -p gpurun_out
timeout 90 python -m pytest tests -x -q -m gpu 2>&1 | tail -25 | tee gpurun_out/pytest_gpu_n3.log
timeout 30 python scripts/diag_refit.py 2>&1 | tail -4 | tee gpurun_out/diag_refit.txt
