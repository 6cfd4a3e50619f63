#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
NSGPU_GRAPH=1 timeout 300 python tools/quick_gpu.py 1024 4096 2>&1 | cut -c1-150
python tools/prof_step.py 4096 2 1 > gpurun_out/c8_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,launch__registers_per_thread,smsp__inst_executed.sum --clock-control none -s 700 -c 450 --csv --log-file gpurun_out/c8_launches.csv python tools/prof_step.py 4096 2 1 > gpurun_out/c8_ncu.log 2>&1
cat gpurun_out/c8_plain.log
