#!/bin/bash
# dataflow kernel with its own per-lane copy cap (16): whole GPU suite, foreign frames, Tier-1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider -x 2>&1 | tail -2
for w in loglike textlike; do timeout 300 python tools/foreign_speed.py 1024 $w 2>&1 | tail -1 | grep -o '"cli_[0-9A-Za-z_]*": {"GBps": [0-9.]*' | tr '\n' ' '; echo; done | tee gpurun_out/r3p_df_lanecap16.txt
timeout 600 python tools/tier1_speed.py 1024 2>&1 | grep -v "^\[b2lz4" | grep -v "^reference" | tail -5 | tee -a gpurun_out/r3p_df_lanecap16.txt
