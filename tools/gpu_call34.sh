#!/bin/bash
echo "--- SOR coarsest (auto)"; timeout 300 python tools/aspect_cycles.py 2>&1 | tail -6
echo "--- plain GS coarsest"; NSGPU_COARSE_OMEGA=1 timeout 300 python tools/aspect_cycles.py 2>&1 | tail -6
