#!/bin/bash
mkdir -p gpurun_out
python tools/prof_step.py 4096 2 1 > gpurun_out/c5_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,launch__registers_per_thread,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 700 -c 450 --csv --log-file gpurun_out/c5_launches.csv python tools/prof_step.py 4096 2 1 > gpurun_out/c5_ncu.log 2>&1
cat gpurun_out/c5_plain.log
