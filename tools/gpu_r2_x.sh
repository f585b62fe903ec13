#!/bin/bash
# round 2: compacted second measuring pass of find_emit_kernel (product) against the old per-lane rounds (variant nocompact);
# output-tile variants of the decoder once more (cheaper whole-warp copies)
mkdir -p gpurun_out
for w in loglike textlike; do
  echo "==== $w"; BENCH_ARGS="--workload $w" bash tools/gpu_variants.sh 2>&1 | grep -v "^$"; cp gpurun_out/variants.log gpurun_out/r2x_variants_$w.log
done
echo "== parity (encoder) with the product build"
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_lz4f.py -q -m gpu -p no:cacheprovider 2>&1 | tail -3
echo "== ncu find_emit"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:find_emit_kernel -s 3 -c 1 -f -o gpurun_out/r2x_full_find_emit_kernel \
   python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded > gpurun_out/ncu_fe.log 2>&1; tail -1 gpurun_out/ncu_fe.log
