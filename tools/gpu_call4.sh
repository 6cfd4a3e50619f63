#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
echo "== direct"; timeout 300 python tools/quick_gpu.py 1024 4096 2>&1 | tail -4
echo "== graph"; NSGPU_GRAPH=1 timeout 300 python tools/quick_gpu.py 1024 4096 2>&1 | tail -6
for pf in 2 4 6; do echo "== PF=$pf"; NSGPU_PF=$pf timeout 300 python tools/quick_gpu.py 4096 2>&1 | head -1; done
echo "== graph pytest"; NSGPU_GRAPH=1 timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
