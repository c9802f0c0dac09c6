#!/bin/bash
mkdir -p gpurun_out
timeout 100 python -m pytest tests -x -q -m gpu 2>&1 | tail -25 | tee gpurun_out/pytest_gpu_n2b.log
timeout 40 python scripts/diag_projnormal.py 2>&1 | tail -6 | tee gpurun_out/diag_projnormal.txt
