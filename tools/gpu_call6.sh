#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
echo "== direct"; timeout 300 python tools/quick_gpu.py 1024 4096 2>&1 | tail -4
echo "== graph"; NSGPU_GRAPH=1 timeout 300 python tools/quick_gpu.py 4096 2>&1 | tail -2
for pf in 2 4 6; do echo "== PF=$pf"; NSGPU_PF=$pf timeout 300 python tools/quick_gpu.py 4096 2>&1 | head -1; done
python tools/prof_step.py 4096 2 1 > gpurun_out/c6_plain.log 2>&1 && \
ncu --set full --import-source on --clock-control none -k regex:k_march -s 184 -c 2 -o gpurun_out/c6_march_fused python tools/prof_step.py 4096 2 1 > gpurun_out/c6_ncu.log 2>&1
cat gpurun_out/c6_plain.log; tail -2 gpurun_out/c6_ncu.log
