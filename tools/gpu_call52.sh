#!/bin/bash
mkdir -p gpurun_out
for v in A B; do
cp build/ab/$v.so navierstokessolver_b200/libnsgpu.so
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/c52_$v.json; python - $v <<'PY'
import json,sys; d=json.load(open("gpurun_out/c52_%s.json"%sys.argv[1])); print(sys.argv[1], d["ms_per_step"], d["phase_ms_last_step"], d["roofline"]["achieved"])
PY
done
NSGPU_TRACE=1 NSGPU_GRAPH=0 timeout 300 python bench.py --graph 0 --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/c52_trace.txt | head -c 10; grep "nsgpu trace" gpurun_out/c52_trace.txt
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
