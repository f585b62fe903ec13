#!/bin/bash
# round 2: pooled engine contexts + larger pinned pool in the LZ4F layer — LZ4F / conduit / CLI tests, Tier-1 speeds twice
mkdir -p gpurun_out
echo "== pytest"; timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider 2>&1 | tail -4
for i in 1 2; do
echo "== tier1 run $i"; B2LZ4_TRACE=1 timeout 600 python tools/tier1_speed.py 1024 > gpurun_out/r3h_tier1_$i.txt 2>&1; grep -v "^\[b2lz4" gpurun_out/r3h_tier1_$i.txt | tail -9
done
echo "== tier1 without the context pool"; B2LZ4_CTX_POOL=0 timeout 600 python tools/tier1_speed.py 1024 2>&1 | grep -v "^\[b2lz4" | tail -9 | tee gpurun_out/r3h_tier1_nopool.txt
