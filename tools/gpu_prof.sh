#!/bin/bash
# usage: gpu_prof.sh <kernel-regex> <out-name> [size-mib]   — plain run first, then one ncu --set full capture
K=$1; O=$2; M=${3:-256}
python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e --size-mib $M ${BENCH_ARGS} > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$K -s 2 -c 1 -f -o gpurun_out/$O python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e --size-mib $M ${BENCH_ARGS} > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
