#!/bin/bash
# round 2: the driver's 8-GPU launch of bench.py (torchrun, one rank per GPU) — `sharded` block at n_devices = 8 included
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r3x_bench_8gpu.json 2> gpurun_out/r3x_bench_8gpu.err
echo "rc=$?"; tail -c 2500 gpurun_out/r3x_bench_8gpu.json; tail -5 gpurun_out/r3x_bench_8gpu.err
