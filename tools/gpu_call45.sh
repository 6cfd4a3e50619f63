#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c45_bench1.json; python - <<'PY'
import json; d=json.load(open("gpurun_out/c45_bench1.json")); print(d["ms_per_step"], d["phase_ms_last_step"], d["solver_iterations_last_step"], d["residuals_last_step"])
PY
NSGPU_RHS_MARCH=0 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c45_bench1_old.json; python - <<'PY'
import json; d=json.load(open("gpurun_out/c45_bench1_old.json")); print(d["ms_per_step"], d["phase_ms_last_step"], d["solver_iterations_last_step"], d["residuals_last_step"])
PY
