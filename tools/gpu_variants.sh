#!/bin/bash
# bench the product build and every tuning build under lz4-frame-conduit_b200/build/variants (device-resident only)
mkdir -p gpurun_out
for f in lz4-frame-conduit_b200/libb200lz4.so lz4-frame-conduit_b200/build/variants/*.so; do
  echo "== $f ${BENCH_ARGS}"
  B2LZ4_LIB=$PWD/$f timeout 300 python bench.py --steps 5 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded ${BENCH_ARGS} 2>&1 | tail -1 | grep -o '"ratio": [0-9.]*\|"decompress_GBps": [0-9.]*\|"compress_GBps": [0-9.]*\|"value": [0-9.]*\|Error.*' | tr '\n' ' '; echo
done 2>&1 | tee gpurun_out/variants.log
