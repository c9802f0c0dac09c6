#!/bin/bash
mkdir -p gpurun_out
VAR_NAME=B2FWN_DOCC VARS="0 7 8" WLS="c3_band c3_full c5_band" bash scripts/gpu_variants.sh
python scripts/diag_queries.py 2>&1 | tail -4
