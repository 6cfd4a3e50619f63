#!/bin/bash
mkdir -p gpurun_out
for f in 1 0 1 0; do
NSGPU_FUSE_BN2=$f timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c61_$f.json; python - $f <<'PY'
import json,sys; d=json.load(open("gpurun_out/c61_%s.json"%sys.argv[1])); print("fuse",sys.argv[1], d["ms_per_step"], d["phase_ms_last_step"])
PY
done
