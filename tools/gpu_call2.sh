#!/bin/bash
mkdir -p gpurun_out
python tools/prof_step.py 4096 2 1 > gpurun_out/c2_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 2004 -c 1100 --csv --log-file gpurun_out/c2_launches.csv python tools/prof_step.py 4096 2 1 > gpurun_out/c2_ncu.log 2>&1
cat gpurun_out/c2_plain.log; tail -3 gpurun_out/c2_ncu.log
python tools/ncu_summary.py gpurun_out/c2_launches.csv > gpurun_out/c2_summary.json; head -c 3000 gpurun_out/c2_summary.json
