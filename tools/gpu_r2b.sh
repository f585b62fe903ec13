#!/bin/bash
# round 2, call b: dataflow decoder on the GPU — full parity suite, foreign-frame speeds with and without it, bench regression
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2b_pytest.log; cat gpurun_out/r2b_pytest.log
timeout 600 python tools/foreign_speed.py 1024 > gpurun_out/r2b_foreign_df1.txt 2>&1; tail -12 gpurun_out/r2b_foreign_df1.txt
B2LZ4_DF=0 timeout 600 python tools/foreign_speed.py 1024 > gpurun_out/r2b_foreign_df0.txt 2>&1; tail -3 gpurun_out/r2b_foreign_df0.txt
timeout 600 python tools/foreign_speed.py 1024 textlike > gpurun_out/r2b_foreign_text.txt 2>&1; tail -2 gpurun_out/r2b_foreign_text.txt
timeout 300 python tools/foreign_speed.py 1024 zeros > gpurun_out/r2b_foreign_zeros.txt 2>&1; tail -2 gpurun_out/r2b_foreign_zeros.txt
timeout 300 python tools/foreign_speed.py 1024 random > gpurun_out/r2b_foreign_random.txt 2>&1; tail -2 gpurun_out/r2b_foreign_random.txt
python bench.py --steps 5 --warmup 3 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; tail -c 1200 gpurun_out/r2b_bench.json
B2LZ4_TRACE=1 timeout 600 python tools/tier1_speed.py 1024 > gpurun_out/r2b_tier1.txt 2>&1; grep -v "^\[b2lz4" gpurun_out/r2b_tier1.txt | tail -8
