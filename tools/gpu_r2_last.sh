#!/bin/bash
# what the driver runs at the end of the round, on the final tree: smoke, GPU suite (-x), the bench line and the reference arm with default flags
mkdir -p gpurun_out



echo "== bench"; timeout 900 python bench.py > gpurun_out/r3y_bench.json 2> gpurun_out/r3y_bench.err; tail -2 gpurun_out/r3y_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3y_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('metric','value','unit','n_gpus','steps','warmup','ms_per_step','dtype','gpu_launches')})
print('e2e',d['e2e']['value'],'roofline',d['roofline']['kernel'],round(d['roofline']['frac'],4),'other',round(d['roofline_other']['frac'],4),'cpu',d['cpu_baseline']['value'],'clocks',d['clocks'])
print('foreign',d['foreign']['decompress_GBps'], {k:(round(v['compress_GBps'],1),round(v['decompress_GBps'],1),v['bit_exact']) for k,v in d['sharded'].items() if isinstance(v,dict) and 'compress_GBps' in v})
PY
