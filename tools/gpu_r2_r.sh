#!/bin/bash
# round 2: dataflow decoder with the unit's recent output in a shared-memory data window (-DB2_DF_SMEMWIN=1) against the product
mkdir -p gpurun_out
V=$PWD/lz4-frame-conduit_b200/build/variants/smemwin.so
echo "== parity with the variant"; B2LZ4_LIB=$V timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider -x 2>&1 | tail -4
for lib in $PWD/lz4-frame-conduit_b200/libb200lz4.so $V; do
  echo "== foreign $(basename $lib)"; B2LZ4_LIB=$lib timeout 300 python tools/foreign_speed.py 1024 2>&1 | tail -1 | grep -o '"cli_[0-9A-Za-z_]*": {"GBps": [0-9.]*' | tr '\n' ' '; echo
  B2LZ4_LIB=$lib timeout 300 python tools/foreign_speed.py 1024 textlike 2>&1 | tail -1 | grep -o '"cli_[0-9A-Za-z_]*": {"GBps": [0-9.]*' | tr '\n' ' '; echo
done 2>&1 | tee gpurun_out/r3k_smemwin_foreign.txt
