#!/bin/bash
# round 2: device-resident speeds over the preference matrix (block size x index x linked x checksums), 1 GiB log-like
mkdir -p gpurun_out
for a in "--block-id 4" "--block-id 4 --no-index" "--block-id 5 --no-index" "--block-id 6" "--block-id 6 --no-index" "--block-id 7 --no-index" \
         "--block-id 4 --linked --no-index" "--block-id 7 --linked --no-index" "--content-checksum" "--block-id 5 --no-index --content-checksum --block-checksum"; do
  printf "%-70s" "$a"; timeout 400 python bench.py --steps 3 --warmup 2 --skip-cpu --skip-e2e --skip-foreign --skip-sharded $a 2>&1 | tail -1 | grep -o '"ratio": [0-9.]*\|"decompress_GBps": [0-9.]*\|"compress_GBps": [0-9.]*\|Error.*\|error.*' | tr '\n' ' '; echo
done | tee gpurun_out/r3e_prefs_matrix.txt
