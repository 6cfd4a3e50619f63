#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/c14_bench2.json 2> gpurun_out/c14_bench2.err
tail -5 gpurun_out/c14_bench2.err; cat gpurun_out/c14_bench2.json | cut -c1-1800
