#!/bin/bash
# ncu of the two single-CTA kernels around the indexed decode
mkdir -p gpurun_out
for k in frame_scan_kernel decode_finalize_kernel; do
timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f -o gpurun_out/r3n_full_$k python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded > gpurun_out/ncu_$k.log 2>&1; tail -1 gpurun_out/ncu_$k.log
done
