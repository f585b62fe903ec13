#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider -x 2>&1 | tail -2
python bench.py --steps 2 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r3o_launches.csv python bench.py --steps 2 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded > gpurun_out/ncu_l.log 2>&1
grep -E "frame_scan|decode_finalize" gpurun_out/r3o_launches.csv | awk -F'","' '{print substr($5,1,30), $NF}' | tail -4
for w in zeros random loglike; do python bench.py --steps 5 --warmup 3 --skip-cpu --skip-e2e --skip-foreign --skip-sharded --workload $w 2>&1 | tail -1 | grep -o '"decompress_GBps": [0-9.]*' ; done
