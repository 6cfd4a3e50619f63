#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for w in 8 16 24 48; do echo "== WPS=$w"; NSGPU_GRAPH=1 NSGPU_WARPS_PER_SM=$w timeout 300 python tools/quick_gpu.py 1024 4096 2>&1 | grep ms_per_step | cut -c1-120; done
python tools/prof_step.py 4096 2 1 > gpurun_out/c7_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,launch__registers_per_thread,smsp__inst_executed.sum --clock-control none -s 700 -c 450 --csv --log-file gpurun_out/c7_launches.csv python tools/prof_step.py 4096 2 1 > gpurun_out/c7_ncu.log 2>&1
cat gpurun_out/c7_plain.log
