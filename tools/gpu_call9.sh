#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
echo "== tail on, graph"; NSGPU_GRAPH=1 timeout 300 python tools/quick_gpu.py 1024 4096 2>&1 | grep ms_per_step | cut -c1-200
echo "== tail off, graph"; NSGPU_TAIL=0 NSGPU_GRAPH=1 timeout 300 python tools/quick_gpu.py 1024 4096 2>&1 | grep ms_per_step | cut -c1-100
echo "== tail on, direct"; timeout 300 python tools/quick_gpu.py 4096 2>&1 | grep ms_per_step | cut -c1-100
echo "== big"; NSGPU_GRAPH=1 timeout 300 python tools/quick_gpu.py 8192 16384 2>&1 | cut -c1-200
