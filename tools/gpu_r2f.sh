#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r2f_pytest.log; cat gpurun_out/r2f_pytest.log
B2LZ4_TRACE=1 timeout 400 python tools/tier1_speed.py 1024 > gpurun_out/r2f_tier1.txt 2>&1; grep -v "^\[b2lz4" gpurun_out/r2f_tier1.txt | tail -8
SUM=1 timeout 200 python tools/config3.py 4 1 | tee gpurun_out/r2f_config3_sum.json
