#!/bin/bash
# First-line GPU check: smoke, parity tests (all, not -x), a short bench.  Everything lands in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -15
echo "== pytest"; timeout 1500 python -m pytest tests -q -m gpu -rA --tb=short -p no:cacheprovider 2>&1 | tail -80 | tee gpurun_out/pytest_gpu.log
echo "== bench"; timeout 600 python bench.py --steps 3 --warmup 3 --size-mib ${BENCH_MIB:-1024} 2>&1 | tail -5 | tee gpurun_out/bench_first.log
