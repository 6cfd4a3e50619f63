#!/bin/bash
mkdir -p gpurun_out
run2() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
for c in "256 128 5" "200 36 3" "130 64 3" "1024 512 3"; do
  run2 29511 tools/mgpu_check.py $c 2>&1 | tail -1 | cut -c1-330
done
NSGPU_RB=0 NSGPU_TRACE=1 NSGPU_GRAPH=0 run2 29512 bench.py --graph 0 --gpus 2 --steps 3 --warmup 3 2> gpurun_out/c38_trace.txt | tail -1 | head -c 10; grep "nsgpu trace rank 0" gpurun_out/c38_trace.txt
b() { run2 29512 bench.py --gpus 2 --steps 10 --warmup 3 2>/dev/null | tail -1 > gpurun_out/c38_$1.json; python - $1 <<'PY'
import json,sys; d=json.load(open("gpurun_out/c38_%s.json"%sys.argv[1])); print(sys.argv[1], d["ms_per_step"], d["phase_ms_last_step"], d["solver_iterations_last_step"], d["roofline"]["achieved"])
PY
}
NSGPU_RB=0 b rb0
b rb12
NSGPU_RB_MINNX=1024 b rb12_min1024
