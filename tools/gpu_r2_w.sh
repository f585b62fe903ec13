#!/bin/bash
# round 2: gather by warp per sub-block + layout with eight loads in flight — whole GPU suite, the four shapes, launch list
mkdir -p gpurun_out
echo "== pytest"; timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider 2>&1 | tail -4
echo "== shapes"; BENCH_ARGS="--skip-foreign --skip-sharded --skip-e2e" bash tools/gpu_shapes.sh | tee gpurun_out/r2w_shapes.txt
echo "== launch list"
python bench.py --steps 2 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2w_launches.csv \
   python bench.py --steps 2 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded > gpurun_out/ncu_l.log 2>&1
grep -E "layout_kernel|gather_kernel|find_emit" gpurun_out/r2w_launches.csv | awk -F'","' '{print $5, $NF}' | sort | uniq -c | head -12
