#!/bin/bash
mkdir -p gpurun_out
run2() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
b() { cp build/ab/$1.so navierstokessolver_b200/libnsgpu.so; run2 29512 bench.py --gpus 2 --steps 10 --warmup 3 2>/dev/null | tail -1 > gpurun_out/c41_$2.json; python - $2 <<'PY'
import json,sys; d=json.load(open("gpurun_out/c41_%s.json"%sys.argv[1])); print(sys.argv[1], d["ms_per_step"], d["phase_ms_last_step"], d["roofline"]["achieved"])
PY
}
export NSGPU_RB=0
NSGPU_REP_CELLS=4096 b C_commit9f1 C
NSGPU_REP_CELLS=4096 b D D_rep4k
NSGPU_REP_CELLS=70000 b D D_rep70k
NSGPU_REP_CELLS=262144 b D D_rep262k
NSGPU_REP_CELLS=1100000 b D D_rep1m
NSGPU_REP_CELLS=4200000 b D D_rep4m
for c in "256 128 5" "200 36 3" "130 64 3" "1024 512 3"; do
  run2 29511 tools/mgpu_check.py $c 2>&1 | tail -1 | cut -c1-330
done
