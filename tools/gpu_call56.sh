#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c56.json; python - <<'PY'
import json,sys; d=json.load(open("gpurun_out/c56.json")); print(d["ms_per_step"], d["phase_ms_last_step"], d["solver_iterations_last_step"], d["residuals_last_step"])
PY
NSGPU_TRACE=1 NSGPU_GRAPH=0 timeout 300 python bench.py --graph 0 --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/c56_trace.txt | head -c 10; grep "nsgpu trace.*99999" gpurun_out/c56_trace.txt
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --ny 1024 --nx-per-gpu 1024 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('1024:', d['ms_per_step'], d['phase_ms_last_step'])"
