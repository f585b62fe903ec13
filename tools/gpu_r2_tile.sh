#!/bin/bash
# round 2: output-tile variants of decode_chains_kernel (-DB2_DEC_TILE=..), parity on the GPU with the variant library, all
# four data shapes device-resident; batch-size sweep of the Tier-1 compress legs; foreign frames first / next file.
mkdir -p gpurun_out
V=lz4-frame-conduit_b200/build/variants
echo "== parity with the tile build"
B2LZ4_LIB=$PWD/$V/tile4096.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py tests/test_gpu_lz4f.py -q -m gpu -p no:cacheprovider 2>&1 | tail -4
for w in loglike textlike zeros random; do
  echo "==== $w"; BENCH_ARGS="--workload $w" bash tools/gpu_variants.sh 2>&1 | grep -v "^$"; cp gpurun_out/variants.log gpurun_out/r2y_variants_$w.log
done
echo "== ncu of the tile build"
B2LZ4_LIB=$PWD/$V/tile4096.so timeout 600 ncu --set full --clock-control none --import-source on -k regex:decode_chains_kernel -s 3 -c 1 -f -o gpurun_out/r2y_full_decode_chains_tile4096 \
   python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded > gpurun_out/ncu_tile.log 2>&1; tail -1 gpurun_out/ncu_tile.log
echo "== tier1 batch sweep"
for mb in 8 12 16 24 32; do B2LZ4_BATCH_MB=$mb timeout 300 python tools/tier1_speed.py 1024 --only-compress 2>&1 | grep -v "^\[b2lz4" | tail -2; done | tee gpurun_out/r2y_tier1_batch_sweep.txt
echo "== tier1"
B2LZ4_TRACE=1 timeout 600 python tools/tier1_speed.py 1024 > gpurun_out/r2y_tier1.txt 2>&1; grep -v "^\[b2lz4" gpurun_out/r2y_tier1.txt | tail -9
