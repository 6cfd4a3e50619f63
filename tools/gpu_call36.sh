#!/bin/bash
mkdir -p gpurun_out
run2() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
t() { NSGPU_TRACE=1 NSGPU_GRAPH=0 run2 29512 bench.py --graph 0 --gpus 2 --steps 5 --warmup 3 2> gpurun_out/c36_$1.txt | tail -1 | head -c 200; echo; echo "== $1"; grep "nsgpu trace rank 0 key *\(10\|21\|22\|120\|121\|5020\|5022\)" gpurun_out/c36_$1.txt; }
NSGPU_P2P=0 t nccl
NSGPU_NOPUSH=1 t arena_nopush
NSGPU_RB=0 t push
