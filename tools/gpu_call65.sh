#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c65.json; python - <<'PY'
import json,sys; d=json.load(open("gpurun_out/c65.json")); print(d["ms_per_step"], d["phase_ms_last_step"], d["residuals_last_step"])
PY
