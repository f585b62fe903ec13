#!/bin/bash
# round 2: block checksums by one warp per block — whole GPU suite, config 2 with and without per-block checksums
mkdir -p gpurun_out
echo "== pytest"; timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider 2>&1 | tail -4
for a in "" "--block-checksum" "--block-checksum --block-id 4 --no-index"; do
  echo "== bench $a"; timeout 300 python bench.py --steps 5 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded $a 2>&1 | tail -1 | grep -o '"ratio": [0-9.]*\|"decompress_GBps": [0-9.]*\|"compress_GBps": [0-9.]*\|"value": [0-9.]*\|Error.*' | tr '\n' ' '; echo
  if [ -n "$a" ]; then echo "== bench $a (thread per block, old build)"; B2LZ4_LIB=$PWD/lz4-frame-conduit_b200/build/variants/old_bx.so timeout 300 python bench.py --steps 5 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded $a 2>&1 | tail -1 | grep -o '"ratio": [0-9.]*\|"decompress_GBps": [0-9.]*\|"compress_GBps": [0-9.]*\|"value": [0-9.]*\|Error.*' | tr '\n' ' '; echo; fi
done | tee gpurun_out/r3d_block_checksum.txt
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r3d_bx_launches.csv python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded --block-checksum > /dev/null 2>&1
grep -E "checksums_kernel" gpurun_out/r3d_bx_launches.csv | awk -F'","' '{print $5, $NF}' | cut -c1-60,200- | tail -4
