#!/bin/bash
# ncu --set full captures of the kernels of a 4096^2 time step (direct launches; ncu serialises graph nodes anyway).
# usage: tools/ncu_round2.sh <tag> <kernel regex> [skip] [n]   ->  gpurun_out/<tag>.ncu-rep + raw csv
tag=$1; rx=$2; skip=${3:-2}; n=${4:-4096}
export NSGPU_GRAPH=0
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$rx" -s $skip -c 1 -f -o gpurun_out/$tag python tools/prof_step.py $n 2 1 > gpurun_out/${tag}.log 2>&1
ncu -i gpurun_out/$tag.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
ncu -i gpurun_out/$tag.ncu-rep --page source --csv > gpurun_out/${tag}_source.csv 2>/dev/null
ls -la gpurun_out/$tag.ncu-rep
