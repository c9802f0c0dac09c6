#!/bin/bash
mkdir -p gpurun_out
VAR_NAME=B2FWN_LINEAR VARS="0" WLS="c1 c3_band c3_full c5_band" bash scripts/gpu_variants.sh
python scripts/diag_e2e.py c3_band 2>&1 | tail -2
