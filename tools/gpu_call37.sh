#!/bin/bash
mkdir -p gpurun_out
run2() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
t() { NSGPU_RB=0 NSGPU_TRACE=1 NSGPU_GRAPH=0 run2 29512 bench.py --graph 0 --gpus 2 --steps 3 --warmup 3 2> gpurun_out/c37_$1.txt | tail -1 | head -c 10; echo "== $1"; grep "nsgpu trace rank [01] key *\(10\|21\|5020\|5022\)" gpurun_out/c37_$1.txt; }
NSGPU_DBG=0 t full
NSGPU_DBG=1 t nowait
NSGPU_DBG=2 t nodata
NSGPU_DBG=5 t nowait_nosignal
NSGPU_DBG=7 t none
