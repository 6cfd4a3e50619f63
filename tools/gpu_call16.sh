#!/bin/bash
mkdir -p gpurun_out
R="timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
$R tools/mgpu_check.py 256 128 5 > gpurun_out/c16_a.log 2>&1; grep '"world"' gpurun_out/c16_a.log || tail -5 gpurun_out/c16_a.log
$R tools/mgpu_check.py 1024 512 3 > gpurun_out/c16_b.log 2>&1; grep '"world"' gpurun_out/c16_b.log || tail -5 gpurun_out/c16_b.log
$R tools/mgpu_check.py 200 36 3 > gpurun_out/c16_c.log 2>&1; grep '"world"' gpurun_out/c16_c.log || tail -5 gpurun_out/c16_c.log
$R tools/mgpu_check.py 4096 4096 3 > gpurun_out/c16_d.log 2>&1; grep '"world"' gpurun_out/c16_d.log || tail -5 gpurun_out/c16_d.log
