#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c60.json; python - <<'PY'
import json,sys; d=json.load(open("gpurun_out/c60.json")); print(d["ms_per_step"], d["phase_ms_last_step"], d["solver_iterations_last_step"], d["residuals_last_step"], d["gpu_launches"])
PY
run2() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
for c in "256 128 5" "130 64 3"; do
  run2 29511 tools/mgpu_check.py $c 2>&1 | tail -1 | cut -c1-330
done
run2 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/c60_bench2.json; python - <<'PY'
import json; d=json.load(open("gpurun_out/c60_bench2.json")); print(d["ms_per_step"], d["phase_ms_last_step"], d["solver_iterations_last_step"])
PY
