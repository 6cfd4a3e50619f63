#!/bin/bash
# ncu --set full captures of the kernels a 4096^2 time step spends most of its time in besides the plain sweep:
# the fused fine-level passes of the phi V-cycle (down: 2 sweeps + residual restriction; up: prolongation + 2 sweeps + norm)
# and the projection.  Direct launches (ncu serialises graph nodes anyway).  Output: gpurun_out/<tag>_*.ncu-rep
tag=${1:-r1_h}
export NSGPU_GRAPH=0
common="--set full --clock-control none --import-source on --kernel-name-base demangled"
ncu $common -k 'regex:k_march<\(int\)2, \(bool\)1, \(int\)2' -s 4 -c 1 -f -o gpurun_out/${tag}_fused_up python tools/prof_step.py 4096 2 1 > gpurun_out/${tag}_ncu_up.log 2>&1
ncu $common -k 'regex:k_march<\(int\)2, \(bool\)0, \(int\)1' -s 0 -c 1 -f -o gpurun_out/${tag}_fused_down python tools/prof_step.py 4096 2 1 > gpurun_out/${tag}_ncu_down.log 2>&1
[ -n "$SKIP_CORRECT" ] || ncu $common -k 'regex:k_correct' -s 2 -c 1 -f -o gpurun_out/${tag}_correct python tools/prof_step.py 4096 2 1 > gpurun_out/${tag}_ncu_correct.log 2>&1
ls -la gpurun_out/*.ncu-rep
