#!/bin/bash
mkdir -p gpurun_out
NSGPU_TAIL_DBG=1 python tools/prof_step.py 4096 2 1 2>&1 | tail -30
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c59.json; python - <<'PY'
import json,sys; d=json.load(open("gpurun_out/c59.json")); print(d["ms_per_step"], d["phase_ms_last_step"], d["solver_iterations_last_step"], d["residuals_last_step"])
PY
