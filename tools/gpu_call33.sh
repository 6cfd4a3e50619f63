#!/bin/bash
mkdir -p gpurun_out
run2() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
for c in "256 128 5" "200 36 3" "130 64 3" "1024 512 3"; do
  run2 29511 tools/mgpu_check.py $c 2>&1 | tail -1 | cut -c1-330
done
NSGPU_TRACE=1 NSGPU_GRAPH=0 run2 29512 bench.py --graph 0 --gpus 2 --steps 5 --warmup 3 2> gpurun_out/c33_trace2.txt | tail -1 | head -c 300
echo
grep "nsgpu trace rank 0" gpurun_out/c33_trace2.txt
echo "--- bench 2 ranks (graphs)"
run2 29512 bench.py --gpus 2 --steps 10 --warmup 3 2>gpurun_out/c33_b2.err | tail -1 > gpurun_out/c33_bench2.json; python - <<'PY'
import json; d=json.load(open("gpurun_out/c33_bench2.json")); print(d["ms_per_step"], d["phase_ms_last_step"], d["solver_iterations_last_step"])
PY
NSGPU_RB=0 run2 29512 bench.py --gpus 2 --steps 10 --warmup 3 2>/dev/null | tail -1 > gpurun_out/c33_bench2_rb0.json; python - <<'PY'
import json; d=json.load(open("gpurun_out/c33_bench2_rb0.json")); print("rb0", d["ms_per_step"], d["phase_ms_last_step"])
PY
