#!/bin/bash
mkdir -p gpurun_out
python tools/prof_step.py 4096 2 1 > gpurun_out/c11_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,launch__registers_per_thread,smsp__inst_executed.sum --clock-control none -s 400 -c 300 --csv --log-file gpurun_out/c11_launches.csv python tools/prof_step.py 4096 2 1 > gpurun_out/c11_ncu.log 2>&1
cat gpurun_out/c11_plain.log
