#!/bin/bash
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
echo "== auto"; NSGPU_GRAPH=1 timeout 600 python tools/quick_gpu.py 1024 4096 16384 2>&1 | grep ms_per_step | cut -c1-250
echo "== mg"; NSGPU_VEL_SOLVER=mg NSGPU_GRAPH=1 timeout 600 python tools/quick_gpu.py 4096 2>&1 | grep ms_per_step | cut -c1-120
echo "== sor"; NSGPU_VEL_SOLVER=sor NSGPU_GRAPH=1 timeout 600 python tools/quick_gpu.py 4096 16384 2>&1 | grep ms_per_step | cut -c1-120
echo "== graph pytest"; NSGPU_GRAPH=1 timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
