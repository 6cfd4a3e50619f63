#!/bin/bash
mkdir -p gpurun_out
for pfd in 0 4 8 12 16 24; do
NSGPU_PFD=$pfd timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c51_$pfd.json; python - $pfd <<'PY'
import json,sys; d=json.load(open("gpurun_out/c51_%s.json"%sys.argv[1])); print("pfd",sys.argv[1], d["ms_per_step"], d["phase_ms_last_step"]["velocity_solves"], d["phase_ms_last_step"]["phi_solve"], d["roofline"]["achieved"])
PY
done
