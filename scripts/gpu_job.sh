#!/bin/bash
mkdir -p gpurun_out
timeout 150 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 | tee gpurun_out/pytest_gpu_n2.log
timeout 60 python scripts/diag_sign.py 2>&1 | tail -6 | tee gpurun_out/diag_sign.txt
