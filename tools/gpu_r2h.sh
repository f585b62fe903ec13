#!/bin/bash
# round 2, call h: back-off of waiting warps in the dataflow decoder (default 64..1024 ns, and two variants)
mkdir -p gpurun_out
timeout 200 python tools/foreign_speed.py 1024 > gpurun_out/r2h_foreign_default.txt 2>&1; tail -1 gpurun_out/r2h_foreign_default.txt | cut -c1-1600
for v in nap256_4096 nap32_256; do
  B2LZ4_LIB=$PWD/lz4-frame-conduit_b200/build/variants/$v.so timeout 200 python tools/foreign_speed.py 1024 > gpurun_out/r2h_foreign_$v.txt 2>&1; echo $v; tail -1 gpurun_out/r2h_foreign_$v.txt | cut -c1-1600
done
B2LZ4_DF_WARPS=8 timeout 200 python tools/foreign_speed.py 1024 > gpurun_out/r2h_foreign_w8.txt 2>&1; echo w8; tail -1 gpurun_out/r2h_foreign_w8.txt | cut -c1-1000
