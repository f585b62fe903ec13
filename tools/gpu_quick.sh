#!/bin/bash
# Quick GPU iteration: parity tests then a short device-resident bench.  Output in gpurun_out/.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --tb=short -p no:cacheprovider 2>&1 | tail -15 | tee gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --skip-cpu ${BENCH_ARGS} 2>&1 | tail -3 | tee gpurun_out/bench_quick.log
