#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>gpurun_out/c63.err > gpurun_out/c63.json; tail -3 gpurun_out/c63.err; python - <<'PY'
import json,sys; d=json.load(open("gpurun_out/c63.json")); print(d["ms_per_step"], d["e2e"], d["phase_ms_last_step"])
PY
