#!/bin/bash
mkdir -p gpurun_out
for v in P2 P3 P4 P6; do
cp build/ab/$v.so navierstokessolver_b200/libnsgpu.so
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c53_$v.json; python - $v <<'PY'
import json,sys; d=json.load(open("gpurun_out/c53_%s.json"%sys.argv[1])); print(sys.argv[1], d["ms_per_step"], d["phase_ms_last_step"]["velocity_solves"], d["phase_ms_last_step"]["phi_solve"], d["roofline"]["achieved"])
PY
timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --ny 16384 --nx-per-gpu 16384 2>/dev/null > gpurun_out/c53_${v}_16k.json; python - $v <<'PY'
import json,sys; d=json.load(open("gpurun_out/c53_%s_16k.json"%sys.argv[1])); print(sys.argv[1], "16k", d["ms_per_step"], d["phase_ms_last_step"]["velocity_solves"], d["phase_ms_last_step"]["phi_solve"])
PY
done
