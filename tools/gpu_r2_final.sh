#!/bin/bash
# round 2, final 1-GPU pass (second edition, after the layout / gather rewrite and the 12 MiB compress batches): smoke, full parity suite, bench (all legs), reference arm, shapes, Tier-1 speeds, foreign frames,
# ncu launch lists and --set full captures of the four codec kernels.  Everything lands in gpurun_out/r3z_*.
mkdir -p gpurun_out
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
echo "== pytest"; timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider 2>&1 | tail -12 | tee gpurun_out/r3z_pytest.log
echo "== bench"; timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r3z_bench.json 2> gpurun_out/r3z_bench.err; tail -c 400 gpurun_out/r3z_bench.json; tail -2 gpurun_out/r3z_bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r3z_bench_reference.json 2> gpurun_out/r3z_bench_reference.err; tail -c 300 gpurun_out/r3z_bench_reference.json
echo "== shapes"; BENCH_ARGS="--skip-foreign --skip-sharded" bash tools/gpu_shapes.sh | tee gpurun_out/r3z_shapes.txt
echo "== tier1"; B2LZ4_TRACE=1 timeout 500 python tools/tier1_speed.py 1024 > gpurun_out/r3z_tier1.txt 2>&1; grep -v "^\[b2lz4" gpurun_out/r3z_tier1.txt | tail -9
echo "== foreign"; timeout 200 python tools/foreign_speed.py 1024 > gpurun_out/r3z_foreign_log.txt 2>&1; tail -1 gpurun_out/r3z_foreign_log.txt | cut -c1-600
echo "== launch lists"
python bench.py --steps 2 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r3z_launches.csv \
   python bench.py --steps 2 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded > gpurun_out/ncu_l.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r3z_foreign_launches.csv python tools/foreign_speed.py 256 > gpurun_out/ncu_l2.log 2>&1
echo "== ncu full"
for k in decode_chains_kernel find_emit_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f -o gpurun_out/r3z_full_$k \
     python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded > gpurun_out/ncu_$k.log 2>&1; tail -1 gpurun_out/ncu_$k.log
done
timeout 400 ncu --set full --clock-control none --import-source on -k regex:decode_df_kernel -s 2 -c 1 -f -o gpurun_out/r3z_full_decode_df_kernel python tools/foreign_speed.py 256 > gpurun_out/ncu_df.log 2>&1; tail -1 gpurun_out/ncu_df.log
timeout 400 ncu --set full --clock-control none --import-source on -k regex:df_links_kernel -s 2 -c 1 -f -o gpurun_out/r3z_full_df_links_kernel python tools/foreign_speed.py 256 > gpurun_out/ncu_dfl.log 2>&1; tail -1 gpurun_out/ncu_dfl.log
ls -la gpurun_out/*.ncu-rep
