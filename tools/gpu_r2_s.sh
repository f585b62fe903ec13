#!/bin/bash
# round 2: few-block calls through the dataflow kernels (SCAN_DF_FEW) — whole suite, Tier-1 speeds, the full bench line, and both with B2LZ4_DF_FEW=0
mkdir -p gpurun_out
echo "== pytest"; timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider 2>&1 | tail -4
echo "== tier1"; B2LZ4_TRACE=1 timeout 600 python tools/tier1_speed.py 1024 > gpurun_out/r3i_tier1.txt 2>&1; grep -v "^\[b2lz4" gpurun_out/r3i_tier1.txt | grep -v "^reference" | tail -6
echo "== tier1 FEW=0"; B2LZ4_DF_FEW=0 timeout 600 python tools/tier1_speed.py 1024 2>&1 | grep -v "^\[b2lz4" | grep -v "^reference" | tail -6 | tee gpurun_out/r3i_tier1_few0.txt
echo "== bench"; timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r3i_bench.json 2> gpurun_out/r3i_bench.err; tail -2 gpurun_out/r3i_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3i_bench.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'foreign',d['foreign']['decompress_GBps'],d['foreign']['e2e_decompress_GBps'])
print({k:(round(v['compress_GBps'],1),round(v['decompress_GBps'],1),v['bit_exact'],v['frame_xxh32']) for k,v in d['sharded'].items() if isinstance(v,dict) and 'compress_GBps' in v})
PY
echo "== bench FEW=0 (sharded leg)"; B2LZ4_DF_FEW=0 timeout 900 python bench.py --steps 3 --warmup 3 --skip-cpu --skip-foreign > gpurun_out/r3i_bench_few0.json 2>/dev/null
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3i_bench_few0.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'])
print({k:(round(v['compress_GBps'],1),round(v['decompress_GBps'],1)) for k,v in d['sharded'].items() if isinstance(v,dict) and 'compress_GBps' in v})
PY
