#!/bin/bash
mkdir -p gpurun_out
run2() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
b() { cp build/ab/$1.so navierstokessolver_b200/libnsgpu.so; run2 29512 bench.py --gpus 2 --steps 10 --warmup 3 2>/dev/null | tail -1 > gpurun_out/c40_$2.json; python - $2 <<'PY'
import json,sys; d=json.load(open("gpurun_out/c40_%s.json"%sys.argv[1])); print(sys.argv[1], d["ms_per_step"], d["phase_ms_last_step"], d["roofline"]["achieved"])
PY
}
export NSGPU_RB=0
NSGPU_REP_CELLS=4096 b C_commit9f1 C
NSGPU_REP_CELLS=4096 b A_minblk A_rep4k
NSGPU_REP_CELLS=4096 b B_nominblk B_rep4k
NSGPU_REP_CELLS=1100000 b B_nominblk B_rep1m
NSGPU_REP_CELLS=4096 b C_commit9f1 C_again
