#!/bin/bash
# round 2, call e: full GPU suite with the f4 tests, Tier-1 speeds with the smaller foreign batches, bench with all legs (1 GPU)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r2e_pytest.log; cat gpurun_out/r2e_pytest.log
B2LZ4_TRACE=1 timeout 400 python tools/tier1_speed.py 1024 > gpurun_out/r2e_tier1.txt 2>&1; grep -v "^\[b2lz4" gpurun_out/r2e_tier1.txt | tail -8
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err; python - <<'PY'
import json
j=json.loads(open("gpurun_out/r2e_bench.json").read().strip().splitlines()[-1])
print("value", j["value"], "e2e", j["e2e"]["value"]); print(json.dumps(j.get("sharded"))[:1500]); print(json.dumps(j.get("foreign"))[:600])
PY
tail -3 gpurun_out/r2e_bench.err
