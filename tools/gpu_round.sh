#!/bin/bash
# Full GPU pass: smoke + parity tests + bench, then ncu launch list and --set full captures of the two codec kernels.
mkdir -p gpurun_out
bash tools/gpu_check.sh
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; tail -1 gpurun_out/bench_full.json
python bench.py --steps 2 --warmup 3 --skip-cpu --skip-e2e > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 2 --warmup 3 --skip-cpu --skip-e2e > gpurun_out/ncu_l.log 2>&1
for k in decode_chains_kernel find_emit_kernel; do
ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f -o gpurun_out/full_$k \
   python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e > gpurun_out/ncu_$k.log 2>&1; tail -1 gpurun_out/ncu_$k.log
done
