#!/bin/bash
mkdir -p gpurun_out
R="timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
for a in "256 128 5" "1024 512 3" "200 36 3" "4096 4096 3"; do
  $R tools/mgpu_check.py $a > gpurun_out/c18.log 2>&1; grep '"world"' gpurun_out/c18.log || tail -8 gpurun_out/c18.log
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline 2> gpurun_out/c18.err | cut -c1-330
tail -2 gpurun_out/c18.err
echo "== P2P off"; NSGPU_P2P=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline 2> gpurun_out/c18b.err | cut -c1-330
