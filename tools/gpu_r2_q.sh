#!/bin/bash
# round 2: per-lane copy variants of the decoders (eight loads in flight; lane cap 16 / 32) on log-like and text-like data, and on foreign frames
mkdir -p gpurun_out
for w in loglike textlike; do
  echo "==== $w"; BENCH_ARGS="--workload $w" bash tools/gpu_variants.sh 2>&1 | grep -v "^$"
done | tee gpurun_out/r3l_copy_variants.txt
for lib in lz4-frame-conduit_b200/libb200lz4.so lz4-frame-conduit_b200/build/variants/*.so; do
  printf "%-60s" $lib; B2LZ4_LIB=$PWD/$lib timeout 300 python tools/foreign_speed.py 1024 2>&1 | tail -1 | grep -o '"cli_[0-9A-Za-z_]*": {"GBps": [0-9.]*' | tr '\n' ' '; echo
done | tee -a gpurun_out/r3l_copy_variants.txt
