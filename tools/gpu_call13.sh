#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/c13_bench.json 2> gpurun_out/c13_bench.err; tail -3 gpurun_out/c13_bench.err; cat gpurun_out/c13_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/c13_bench_ref.json 2>&1; cat gpurun_out/c13_bench_ref.json
python tools/prof_sweep.py 4096 6 > gpurun_out/c13_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_march -s 2 -c 2 -o gpurun_out/r1_c_sweep_march python tools/prof_sweep.py 4096 6 > gpurun_out/c13_ncu.log 2>&1
cat gpurun_out/c13_plain.log
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --graph 0 > gpurun_out/c13_plain2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r1_c_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --graph 0 > gpurun_out/c13_ncu2.log 2>&1
tail -2 gpurun_out/c13_ncu2.log | cut -c1-300
