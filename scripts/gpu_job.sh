#!/bin/bash
mkdir -p gpurun_out
timeout 70 python -m pytest tests -x -q -m gpu 2>&1 | tail -25 | tee gpurun_out/pytest_gpu_final.log
timeout 25 python scripts/diag_projbricks.py 2>&1 | tail -30 | tee gpurun_out/diag_projbricks.txt
