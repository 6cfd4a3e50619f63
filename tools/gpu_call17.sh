#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 900 python tools/tolerance_gap.py > gpurun_out/c17_tolerance_gap.json 2> gpurun_out/c17_tol.err; tail -3 gpurun_out/c17_tol.err; head -c 600 gpurun_out/c17_tolerance_gap.json
