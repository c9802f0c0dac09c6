#!/bin/bash
mkdir -p gpurun_out
VAR_NAME=B2FWN_SPLIT VARS="1 0" WLS="c1 c3_band c3_full c5_band" bash scripts/gpu_variants.sh
