#!/bin/bash
mkdir -p gpurun_out
for v in A B A B; do
cp build/ab/$v.so navierstokessolver_b200/libnsgpu.so
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c66_$v.json; python - $v <<'PY'
import json,sys; d=json.load(open("gpurun_out/c66_%s.json"%sys.argv[1])); print(sys.argv[1], d["ms_per_step"], d["phase_ms_last_step"]["div_rhs"], d["phase_ms_last_step"]["correct"], d["phase_ms_last_step"]["phi_solve"])
PY
done
