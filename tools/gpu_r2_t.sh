#!/bin/bash
# round 2: plain frames of small blocks with and without the walker's prefetching companion; foreign frames; whole suite
mkdir -p gpurun_out

for a in "--block-id 4 --no-index" "--block-id 5 --no-index" "--block-id 6 --no-index"; do
  for lib in lz4-frame-conduit_b200/libb200lz4.so lz4-frame-conduit_b200/build/variants/*.so; do
  printf "%-28s %-55s" "$a" "$lib"; B2LZ4_LIB=$PWD/$lib timeout 400 python bench.py --steps 5 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded $a 2>&1 | tail -1 | grep -o '"decompress_GBps": [0-9.]*\|"compress_GBps": [0-9.]*\|Error.*\|error.*' | tr '\n' ' '; echo
  done
done | tee gpurun_out/r3q_walk_nap_sweep.txt
