#!/bin/bash
# per-pass timing (NSGPU_TRACE, direct launches): one rank against two
mkdir -p gpurun_out
NSGPU_TRACE=1 NSGPU_GRAPH=0 timeout 300 python bench.py --graph 0 --steps 5 --warmup 3 2> gpurun_out/c32_trace1.txt | head -c 300
echo
NSGPU_TRACE=1 NSGPU_GRAPH=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --graph 0 --gpus 2 --steps 5 --warmup 3 2> gpurun_out/c32_trace2.txt | tail -1 | head -c 300
echo
grep "nsgpu trace" gpurun_out/c32_trace1.txt
grep "nsgpu trace rank 0" gpurun_out/c32_trace2.txt
