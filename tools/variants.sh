#!/bin/bash
# Tuning builds: tools/variants.sh NAME "-DB2_X=.. -DB2_Y=.." ...  ->  lz4-frame-conduit_b200/build/variants/NAME.so
# (select at run time with B2LZ4_LIB=...; the product build is csrc/Makefile)
cd "$(dirname "$0")/../lz4-frame-conduit_b200/csrc" || exit 1
mkdir -p ../build/variants
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC $flags -shared -o ../build/variants/$name.so decode.cu decode_df.cu encode.cu xxh32.cu engine.cu lz4f_compat.cu noalloc.cu &
done
wait
ls -la ../build/variants
