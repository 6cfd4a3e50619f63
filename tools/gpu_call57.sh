#!/bin/bash
mkdir -p gpurun_out
export NSGPU_GRAPH=0
python tools/prof_step.py 4096 2 1 > gpurun_out/c57_plain.log 2>&1 && \
ncu --set full --import-source on --sampling-interval 0 --clock-control none -k regex:k_coarse_tail -s 12 -c 1 -o gpurun_out/c57_tail python tools/prof_step.py 4096 2 1 > gpurun_out/c57_ncu.log 2>&1
cat gpurun_out/c57_plain.log; tail -2 gpurun_out/c57_ncu.log
