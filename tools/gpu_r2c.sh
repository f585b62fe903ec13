#!/bin/bash
# round 2, call c: CTA-local dataflow decoder on the GPU.  Guarded: a slow first step ends the call.
mkdir -p gpurun_out
timeout 150 python tools/foreign_speed.py 256 > gpurun_out/r2c_foreign_256.txt 2>&1; echo "rc=$?"; tail -9 gpurun_out/r2c_foreign_256.txt | cut -c1-400
ok=$(python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/r2c_foreign_256.txt").read().strip().splitlines()[-1])
    print(1 if j["cli_B7_indep"]["GBps"] > 5 and j["ref_conduit_linked"]["GBps"] > 0.3 else 0)
except Exception as e:
    print(0)
PY
)
echo "gate=$ok"
if [ "$ok" != "1" ]; then exit 0; fi
timeout 300 python tools/foreign_speed.py 1024 > gpurun_out/r2c_foreign_log.txt 2>&1; tail -1 gpurun_out/r2c_foreign_log.txt | cut -c1-3000
timeout 300 python tools/foreign_speed.py 1024 textlike > gpurun_out/r2c_foreign_text.txt 2>&1; tail -1 gpurun_out/r2c_foreign_text.txt | cut -c1-3000
timeout 700 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2c_pytest.log; cat gpurun_out/r2c_pytest.log
B2LZ4_TRACE=1 timeout 400 python tools/tier1_speed.py 512 > gpurun_out/r2c_tier1.txt 2>&1; grep -v "^\[b2lz4" gpurun_out/r2c_tier1.txt | tail -8
