#!/bin/bash
mkdir -p gpurun_out
for v in L2_4 L1_2 L1_4 L2_4; do
cp build/ab/$v.so navierstokessolver_b200/libnsgpu.so
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c54_$v.json; python - $v <<'PY'
import json,sys; d=json.load(open("gpurun_out/c54_%s.json"%sys.argv[1])); print(sys.argv[1], d["ms_per_step"], d["phase_ms_last_step"]["velocity_solves"], d["phase_ms_last_step"]["phi_solve"], d["roofline"]["achieved"])
PY
done
