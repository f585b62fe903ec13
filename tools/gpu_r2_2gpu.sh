#!/bin/bash
# the driver's 2-GPU launch of bench.py with the current tree; the multi-device GPU tests
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu -p no:cacheprovider 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r3j_bench_2gpu.json 2> gpurun_out/r3j_bench_2gpu.err
echo "rc=$?"; tail -3 gpurun_out/r3j_bench_2gpu.err | cut -c1-200
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3j_bench_2gpu.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'n',d['n_gpus'])
print({k:(round(v['compress_GBps'],1),round(v['decompress_GBps'],1),v['bit_exact'],v['frame_xxh32']) for k,v in d['sharded'].items() if isinstance(v,dict) and 'compress_GBps' in v})
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | tail -1 | cut -c1-300
