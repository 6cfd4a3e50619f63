#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/quick_gpu.py 4096 2>&1 | tail -2
python tools/prof_sweep.py 4096 6 > gpurun_out/c3_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_sweep_march -s 2 -c 2 -o gpurun_out/r1_b_sweep_march python tools/prof_sweep.py 4096 6 > gpurun_out/c3_ncu.log 2>&1
cat gpurun_out/c3_plain.log; tail -2 gpurun_out/c3_ncu.log
python tools/prof_step.py 4096 2 1 > gpurun_out/c3_plain2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 1974 -c 1100 --csv --log-file gpurun_out/c3_launches.csv python tools/prof_step.py 4096 2 1 > gpurun_out/c3_ncu2.log 2>&1
python tools/ncu_summary.py gpurun_out/c3_launches.csv > gpurun_out/c3_summary.json
