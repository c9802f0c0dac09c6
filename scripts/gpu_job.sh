#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
python scripts/diag_e2e.py c3_band 2>&1 | tail -2; python scripts/diag_e2e.py c5_band 2>&1 | tail -3
