#!/bin/bash
# Full GPU pass: smoke + parity tests + bench, the other data shapes, ncu launch list and --set full captures of the codec kernels.
mkdir -p gpurun_out
bash tools/gpu_check.sh
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; tail -1 gpurun_out/bench_full.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; tail -1 gpurun_out/bench_ref.json
bash tools/gpu_shapes.sh | tee gpurun_out/shapes.txt
echo "== plain drop-in frames (256 KB blocks, no index)"; python bench.py --steps 5 --warmup 3 --skip-cpu --block-id 5 --sub-log2 18 --no-index > gpurun_out/bench_plain_b5.json 2>/dev/null; tail -1 gpurun_out/bench_plain_b5.json | cut -c1-600
python tools/xxh_speed.py | tee gpurun_out/xxh.txt
python tools/config3.py 5 1 | tee gpurun_out/config3_1gpu.json
B2LZ4_TRACE=1 timeout 600 python tools/tier1_speed.py 1024 > gpurun_out/tier1.txt 2>&1; grep -v "^\[b2lz4" gpurun_out/tier1.txt | tail -8     # the unmodified conduit's calls on LZ4F_*
timeout 300 python tools/lone_chain.py | tee gpurun_out/lone_chain.txt                                                        # one-warp cases (foreign frames)
timeout 300 python tools/fuzz_roundtrip.py 300 $RANDOM | tee gpurun_out/fuzz.txt
python bench.py --steps 2 --warmup 3 --skip-cpu --skip-e2e > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 2 --warmup 3 --skip-cpu --skip-e2e > gpurun_out/ncu_l.log 2>&1
for k in decode_chains_kernel find_emit_kernel; do
ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f -o gpurun_out/full_$k \
   python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e > gpurun_out/ncu_$k.log 2>&1; tail -1 gpurun_out/ncu_$k.log
done
python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e --workload random --size-mib 512 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:copy_stored_kernel -s 3 -c 1 -f -o gpurun_out/full_copy_stored_kernel \
   python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e --workload random --size-mib 512 > gpurun_out/ncu_copy.log 2>&1; tail -1 gpurun_out/ncu_copy.log
