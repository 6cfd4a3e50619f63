#!/bin/bash
mkdir -p gpurun_out
for v in U1 U2 U4; do
cp build/ab/$v.so navierstokessolver_b200/libnsgpu.so
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c67_$v.json; python - $v <<'PY'
import json,sys; d=json.load(open("gpurun_out/c67_%s.json"%sys.argv[1])); print(sys.argv[1], d["ms_per_step"], d["phase_ms_last_step"]["div_rhs"], d["phase_ms_last_step"]["correct"])
PY
done
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
