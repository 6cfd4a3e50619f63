#!/bin/bash
# verification after the PUSH template split: GPU tests, single-GPU bench, 2-rank parity and bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/c30_bench1.json 2> gpurun_out/c30_bench1.err; cat gpurun_out/c30_bench1.json | head -c 1500
for c in "256 128 5" "200 36 3" "130 64 3"; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py $c 2>&1 | tail -3
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 2>/dev/null | tail -1 | head -c 1500
