#!/bin/bash
# per tuning build: device-resident bench line + datagen ratio against the reference default
for f in lz4-frame-conduit_b200/build/variants/*.so; do
  echo "== $f"
  B2LZ4_LIB=$PWD/$f timeout 300 python bench.py --steps 5 --warmup 3 --skip-cpu --skip-e2e ${BENCH_ARGS} 2>&1 | tail -1 | grep -o '"ratio": [0-9.]*\|"decompress_GBps": [0-9.]*\|"compress_GBps": [0-9.]*\|Error.*' | tr '\n' ' '; echo
  B2LZ4_LIB=$PWD/$f timeout 120 python tools/ratio_shapes.py 2>&1 | tail -1
done 2>&1 | tee gpurun_out/variants_ratio.log
