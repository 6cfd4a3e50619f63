#!/bin/bash
mkdir -p gpurun_out
R="timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
for a in "256 128 5" "1024 512 3" "200 36 3" "130 64 3"; do
  $R tools/mgpu_check.py $a > gpurun_out/c20.log 2>&1; grep '"world"' gpurun_out/c20.log | cut -c1-330 || grep -v "^\*\|OMP\|^$" gpurun_out/c20.log | head -12
done > gpurun_out/c20_checks.log 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/c20_bench2.json 2> gpurun_out/c20.err
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/c20_bench1.json 2>> gpurun_out/c20.err
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --graph 0 > gpurun_out/c20_bench1d.json 2>> gpurun_out/c20.err
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/c20_pytest.log
