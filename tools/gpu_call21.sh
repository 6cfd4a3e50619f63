#!/bin/bash
mkdir -p gpurun_out
R="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$R --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/c21_weak8.json 2> gpurun_out/c21_weak8.err
$R --master-port 29522 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu-baseline --ny 16384 --nx-per-gpu 2048 > gpurun_out/c21_strong8_16384.json 2> gpurun_out/c21_strong8.err
$R --master-port 29523 tools/mgpu_check.py 512 128 3 > gpurun_out/c21_check8.log 2>&1
grep '"world"' gpurun_out/c21_check8.log | cut -c1-400
tail -2 gpurun_out/c21_weak8.err | cut -c1-300
