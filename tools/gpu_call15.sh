#!/bin/bash
mkdir -p gpurun_out
R="timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
$R tools/mgpu_check.py 256 128 5 2>&1 | grep -v "^\*\|OMP_NUM\|^$\|W1" | tail -4
$R tools/mgpu_check.py 1024 512 3 2>&1 | grep -v "^\*\|OMP_NUM\|^$\|W1" | tail -3
$R tools/mgpu_check.py 200 36 3 2>&1 | grep -v "^\*\|OMP_NUM\|^$\|W1" | tail -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/c15.err | cut -c1-700
tail -3 gpurun_out/c15.err
