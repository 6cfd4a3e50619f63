#!/bin/bash
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
NSGPU_GRAPH=1 timeout 600 python tools/quick_gpu.py 1024 2048 4096 8192 16384 2>&1 | cut -c1-190
for pf in 3 4; do echo "== PF=$pf"; NSGPU_PF=$pf timeout 300 python tools/quick_gpu.py 4096 16384 2>&1 | grep sweep_ms; done
