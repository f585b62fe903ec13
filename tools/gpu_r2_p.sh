#!/bin/bash
# round 2: copy_stored_kernel with the next slice's look-ups made ahead; slice sizes 64 / 128 / 256 KB (random = every block stored)
mkdir -p gpurun_out
for w in random zeros; do echo "==== $w"; BENCH_ARGS="--workload $w" bash tools/gpu_variants.sh 2>&1 | grep -v "^$"; done | tee gpurun_out/r3m_copy_stored.txt
echo "== parity"; timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py -q -m gpu -p no:cacheprovider 2>&1 | tail -2
