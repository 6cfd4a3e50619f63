#!/bin/bash
# NCCL-only fallback (no peer mapping) still reproduces the single-GPU run
run2() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
for c in "256 128 5" "130 64 3"; do
  NSGPU_P2P=0 run2 29511 tools/mgpu_check.py $c 2>&1 | tail -1 | cut -c1-330
done
NSGPU_P2P=0 run2 29512 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | cut -c1-200
