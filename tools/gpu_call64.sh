#!/bin/bash
mkdir -p gpurun_out
export NSGPU_GRAPH=0
python tools/prof_step.py 4096 2 1 > gpurun_out/c64_plain.log 2>&1 && \
ncu --set full --import-source on --clock-control none -k regex:'k_div_rhs|k_correct|k_sub_mean' -s 6 -c 3 -o gpurun_out/c64_phases python tools/prof_step.py 4096 2 1 > gpurun_out/c64_ncu.log 2>&1
cat gpurun_out/c64_plain.log; tail -2 gpurun_out/c64_ncu.log
