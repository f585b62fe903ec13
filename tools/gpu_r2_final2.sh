#!/bin/bash
# round 2, last 1-GPU pass: the allocation-free entry point alone in a fresh process (plain and under compute-sanitizer), then
# tools/gpu_r2_final.sh (smoke, the whole GPU suite — where the same test runs after all the others —, bench, ncu)
mkdir -p gpurun_out
echo "== no-alloc test alone"; timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider 2>&1 | tail -5 | tee gpurun_out/r3a_noalloc_alone.log
echo "== no-alloc under memcheck"; timeout 600 compute-sanitizer --tool memcheck --log-file gpurun_out/r3a_noalloc_memcheck.txt python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider 2>&1 | tail -3; tail -3 gpurun_out/r3a_noalloc_memcheck.txt
bash tools/gpu_r2_final.sh
