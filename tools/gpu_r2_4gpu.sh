#!/bin/bash
# the driver's 2-GPU launch of bench.py with the current tree; the multi-device GPU tests
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29527 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/r3w_bench_4gpu.json 2> gpurun_out/r3w_bench_4gpu.err
echo "rc=$?"; tail -3 gpurun_out/r3w_bench_4gpu.err | cut -c1-200
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3w_bench_4gpu.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'n',d['n_gpus'])
print({k:(round(v['compress_GBps'],1),round(v['decompress_GBps'],1),v['bit_exact'],v['frame_xxh32']) for k,v in d['sharded'].items() if isinstance(v,dict) and 'compress_GBps' in v})
PY
