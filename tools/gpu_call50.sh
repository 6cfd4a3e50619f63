#!/bin/bash
# round-1 final measurement set (profiles/r1_e_*)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 > gpurun_out/r1_e_bench_4096.json 2> gpurun_out/c50_bench.err; tail -3 gpurun_out/c50_bench.err; cut -c1-400 gpurun_out/r1_e_bench_4096.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1_e_bench_reference_arm.json 2> gpurun_out/c50_ref.err; cut -c1-600 gpurun_out/r1_e_bench_reference_arm.json
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --ny 1024 --nx-per-gpu 1024 2>/dev/null > gpurun_out/r1_e_bench_1024.json; cut -c1-300 gpurun_out/r1_e_bench_1024.json
python bench.py --steps 4 --warmup 3 --no-cpu-baseline --ny 16384 --nx-per-gpu 16384 2>/dev/null > gpurun_out/r1_e_bench_16384.json; cut -c1-300 gpurun_out/r1_e_bench_16384.json
python tools/prof_sweep.py 4096 6 > gpurun_out/c50_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_march -s 2 -c 2 -o gpurun_out/r1_e_sweep_march python tools/prof_sweep.py 4096 6 > gpurun_out/c50_ncu.log 2>&1
cat gpurun_out/c50_plain.log
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --graph 0 > gpurun_out/c50_plain2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r1_e_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --graph 0 > gpurun_out/c50_ncu2.log 2>&1
tail -2 gpurun_out/c50_ncu2.log | cut -c1-300
