#!/bin/bash
# round 2, call d: dataflow decoder with in-batch source chasing; per-kernel times; bench.py with the foreign + sharded legs (1 GPU)
mkdir -p gpurun_out
timeout 200 python tools/foreign_speed.py 1024 > gpurun_out/r2d_foreign_log.txt 2>&1; tail -1 gpurun_out/r2d_foreign_log.txt | cut -c1-2500
timeout 200 python tools/foreign_speed.py 1024 textlike > gpurun_out/r2d_foreign_text.txt 2>&1; tail -1 gpurun_out/r2d_foreign_text.txt | cut -c1-2500
B2LZ4_DF_WARPS=32 timeout 200 python tools/foreign_speed.py 1024 > gpurun_out/r2d_foreign_log_w32.txt 2>&1; tail -1 gpurun_out/r2d_foreign_log_w32.txt | cut -c1-900
B2LZ4_DF_WARPS=8 timeout 200 python tools/foreign_speed.py 1024 > gpurun_out/r2d_foreign_log_w8.txt 2>&1; tail -1 gpurun_out/r2d_foreign_log_w8.txt | cut -c1-900
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2d_foreign_launches.csv python tools/foreign_speed.py 256 > gpurun_out/r2d_ncu.log 2>&1
grep -c . gpurun_out/r2d_foreign_launches.csv
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; tail -c 3500 gpurun_out/r2d_bench.json; tail -3 gpurun_out/r2d_bench.err
